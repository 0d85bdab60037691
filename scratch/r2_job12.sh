#!/bin/bash
set -u
O=gpurun_out/r2l; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
timeout 900 $TR --master-port 29541 tests/multi_rank_check.py > $O/mr_2.log 2>&1; echo "multi_rank_check rc=$?"; grep -v "$F" $O/mr_2.log | tail -6 | cut -c1-200
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 --stages > $O/bench_msmt17_2.log 2>&1; echo "bench rc=$?"; grep -v "$F" $O/bench_msmt17_2.log | tail -15 | cut -c1-260
