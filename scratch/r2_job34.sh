#!/bin/bash
set -u
O=gpurun_out/r2ag; mkdir -p $O
timeout 120 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x > $O/alone.log 2>&1; echo "alone rc=$?"; tail -4 $O/alone.log | cut -c1-300
timeout 150 python -m pytest tests/test_gpu_stages.py tests/test_abi.py tests/test_zz_no_alloc.py -q -m gpu -x -k "scatter or no_alloc or static or symbol" > $O/after.log 2>&1; echo "after rc=$?"; tail -4 $O/after.log | cut -c1-300
