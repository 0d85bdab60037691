#!/bin/bash
# ncu launch lists (per-launch device time) for profiles/: Duke and MSMT17 shapes, one B200
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_duke.csv $CMD > gpurun_out/ncu1.log 2>&1
CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD2 > gpurun_out/plain_msmt.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_msmt.csv $CMD2 > gpurun_out/ncu3.log 2>&1
tail -n 1 gpurun_out/ncu1.log gpurun_out/ncu3.log
