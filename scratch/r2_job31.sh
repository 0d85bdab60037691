#!/bin/bash
set -u
O=gpurun_out/r2ad; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 tests/multi_rank_check.py > $O/mr_2.log 2>&1; echo "mr rc=$?"; grep -i "streamed\|error\|assert" $O/mr_2.log | head -6 | cut -c1-250
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench_retrieval.py --q 100000 --g 1000000 --rerank --steps 2 > $O/bench_rerank_100k_1M_2gpu.log 2>&1; echo "bench rc=$?"; grep "^{" $O/bench_rerank_100k_1M_2gpu.log | cut -c1-1500
