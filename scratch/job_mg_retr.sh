#!/bin/bash
set -u
mkdir -p gpurun_out/retr
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multi_rank_check.py 2>&1 | tail -4
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench_retrieval.py --steps 2 --warmup 1 --score 2>/dev/null | tail -1 > gpurun_out/retr/full_score_2gpu.log; cut -c1-900 gpurun_out/retr/full_score_2gpu.log
