#!/bin/bash
set -u
mkdir -p gpurun_out/f4
timeout 600 python -m pytest tests/test_gpu_losses.py -x -q 2>&1 | tail -25
timeout 300 python bench_losses.py > gpurun_out/f4/bench_losses.log 2>&1; cat gpurun_out/f4/bench_losses.log | tail -8
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_losses.py -x -q -k "shapes or xbm" 2>&1 | tail -5
