#!/bin/bash
# round 2, job 3 (TWO GPUs): both multi-rank paths against one rank (incl. Duke shape), then the scaling benches
set -u
O=gpurun_out/r2c; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
REID_MR_BIG=1 timeout 900 $TR --master-port 29521 tests/multi_rank_check.py > $O/mr_2.log 2>&1; echo "multi_rank_check rc=$?"
grep -v "^W\|OMP_NUM\|\*\*\*" $O/mr_2.log | tail -24 | cut -c1-300
timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 10 --warmup 3 --stages > $O/bench_msmt17_2.log 2>&1; echo "bench msmt17 rc=$?"
grep -v "^W\|OMP_NUM\|\*\*\*" $O/bench_msmt17_2.log | tail -20 | cut -c1-2500
timeout 600 $TR --master-port 29523 bench.py --gpus 2 --workload duke --steps 20 --warmup 3 --stages > $O/bench_duke_2.log 2>&1; echo "bench duke rc=$?"
grep -v "^W\|OMP_NUM\|\*\*\*" $O/bench_duke_2.log | tail -20 | cut -c1-1500
