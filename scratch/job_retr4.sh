#!/bin/bash
set -u
mkdir -p gpurun_out/retr
timeout 300 python -m pytest tests/test_gpu_retrieval.py -x -q 2>&1 | tail -12
timeout 200 python bench_retrieval.py --steps 2 --warmup 1 --score > gpurun_out/retr/full_score2.log 2>&1; tail -1 gpurun_out/retr/full_score2.log | cut -c1-1500
