#!/bin/bash
set -u
O=gpurun_out/r2f; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_retrieval.py -q -m gpu -x -s > $O/pytest.log 2>&1; echo "pytest rc=$?"; grep "a13\]" $O/pytest.log | cut -c1-300; tail -30 $O/pytest.log | cut -c1-250
