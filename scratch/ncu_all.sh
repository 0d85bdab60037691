#!/bin/bash
# all ncu evidence for profiles/ in one gpurun call (one GPU): launch lists + full captures
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_duke.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dist_tc2_kernel -s 2 -c 1 -o gpurun_out/prof_tc2 $CMD > gpurun_out/ncu2.log 2>&1
$CMD2 > gpurun_out/plain_msmt.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_msmt.csv $CMD2 > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'score_sparse|overlap_sparse|select_push|select_thr|select_merge' -s 5 -c 5 -o gpurun_out/prof_msmt_js $CMD2 > gpurun_out/ncu4.log 2>&1
tail -n 1 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log gpurun_out/ncu4.log
