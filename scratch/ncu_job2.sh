#!/bin/bash
set -u
mkdir -p gpurun_out
CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD2 > gpurun_out/plain_msmt.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'score_sparse|overlap_sparse' -s 2 -c 2 -o gpurun_out/prof_msmt_js $CMD2 > gpurun_out/ncu6.log 2>&1
tail -n 3 gpurun_out/ncu6.log
