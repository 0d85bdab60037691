#!/bin/bash
set -u
mkdir -p gpurun_out/retr
timeout 300 python -m pytest tests/test_gpu_retrieval.py -x -q 2>&1 | tail -15
timeout 120 python bench_retrieval.py --q 20000 --g 200000 --steps 2 --warmup 1 > gpurun_out/retr/small.log 2>&1; tail -2 gpurun_out/retr/small.log | cut -c1-900
timeout 200 python bench_retrieval.py --steps 3 --warmup 1 > gpurun_out/retr/full.log 2>&1; tail -2 gpurun_out/retr/full.log | cut -c1-1500
