#!/bin/bash
set -u
O=gpurun_out/r2g; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_retrieval.py -q -m gpu -x -k streamed_rerank > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
timeout 600 python bench_retrieval.py --q 20000 --g 200000 --rerank --steps 2 --warmup 1 > $O/bench_rerank_20k_200k.log 2>&1; echo "rc=$?"; tail -2 $O/bench_rerank_20k_200k.log | cut -c1-1800
timeout 900 python bench_retrieval.py --rerank --steps 2 --warmup 1 > $O/bench_rerank_100k_1M.log 2>&1; echo "rc=$?"; tail -2 $O/bench_rerank_100k_1M.log | cut -c1-1800
nvidia-smi --query-gpu=memory.used --format=csv,noheader
