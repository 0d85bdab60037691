#!/bin/bash
set -u
mkdir -p gpurun_out/retr
timeout 200 python bench_retrieval.py --steps 2 --warmup 1 --score > gpurun_out/retr/full_score.log 2>&1; tail -1 gpurun_out/retr/full_score.log | cut -c1-1600
timeout 200 python bench_retrieval.py --steps 3 --warmup 1 > gpurun_out/retr/full.log 2>&1; tail -1 gpurun_out/retr/full.log | cut -c1-700
CMD="python bench_retrieval.py --q 20000 --g 200000 --steps 1 --warmup 1 --score"
$CMD > gpurun_out/retr/plain_small.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/retr/launches_retrieval.csv $CMD > gpurun_out/retr/ncu.log 2>&1
tail -1 gpurun_out/retr/ncu.log | cut -c1-300
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
