#!/bin/bash
set -u
O=gpurun_out/r2x; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_retrieval.py tests/test_gpu_losses.py -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log | cut -c1-300
timeout 300 python bench_losses.py > $O/bench_losses.log 2>&1; echo "bench_losses rc=$?"; tail -8 $O/bench_losses.log | cut -c1-300
