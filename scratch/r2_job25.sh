#!/bin/bash
set -u
O=gpurun_out/r2y; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_retrieval.py tests/test_gpu_losses.py -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log | cut -c1-300
timeout 600 python bench_retrieval.py --q 100000 --g 1000000 --score > $O/bench_score_100k_1M.log 2>&1; echo "bench rc=$?"; tail -3 $O/bench_score_100k_1M.log | cut -c1-900
