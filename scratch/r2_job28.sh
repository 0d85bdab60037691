#!/bin/bash
set -u
O=gpurun_out/r2ab; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_retrieval.py -q -m gpu -x -k "rerank" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-300
timeout 600 python bench_retrieval.py --q 100000 --g 1000000 --rerank > $O/bench_rerank_100k_1M.log 2>&1; echo "bench rc=$?"; tail -2 $O/bench_rerank_100k_1M.log | cut -c1-1500
