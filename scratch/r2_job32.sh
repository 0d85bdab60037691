#!/bin/bash
set -u
O=gpurun_out/r2ae; mkdir -p $O
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench_retrieval.py --q 100000 --g 1000000 --rerank --steps 3 > $O/bench_rerank_100k_1M_8gpu.log 2>&1; echo "rerank rc=$?"; grep "^{" $O/bench_rerank_100k_1M_8gpu.log | cut -c1-1400
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29622 bench_retrieval.py --q 100000 --g 1000000 --score --steps 3 > $O/bench_score_100k_1M_8gpu.log 2>&1; echo "score rc=$?"; grep "^{" $O/bench_score_100k_1M_8gpu.log | cut -c1-900
