#!/bin/bash
# ncu evidence for profiles/: launch list + full capture (run under gpurun, one GPU)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_duke.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'dist_tc_kernel|jaccard|select_topk' -s 12 -c 8 -o gpurun_out/prof_duke $CMD > gpurun_out/ncu2.log 2>&1
CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD2 > gpurun_out/plain_msmt.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_msmt.csv $CMD2 > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'jaccard|select_topk' -s 4 -c 4 -o gpurun_out/prof_msmt $CMD2 > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu*.log
ls -la gpurun_out
