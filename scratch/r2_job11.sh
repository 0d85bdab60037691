#!/bin/bash
# round 2, job 11 (ONE GPU): ncu evidence of the final build -- launch lists (MSMT17, Duke) and one --set full capture of
# the dominant kernel and of the HBM-bound ones at the MSMT17 shape.  Staged path (REID-c-entry off via --stages? no: the
# launch list must show the path the bench times, so the default single-call path is profiled).
set -u
O=gpurun_out/r2k; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --settle 0 --no-e2e --no-cpu-baseline"
$CMD > $O/plain_msmt17.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $O/launches_msmt17.csv $CMD > $O/ncu1.log 2>&1; echo "ncu launches msmt17 rc=$?"
$CMD > $O/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'dist_tc2_kernel|score_sparse_kernel|select_push_kernel|overlap_sparse_kernel' -s 12 -c 4 -o $O/full_msmt17 $CMD > $O/ncu2.log 2>&1; echo "ncu full rc=$?"
CMDD="python bench.py --workload duke --steps 2 --warmup 3 --settle 0 --no-e2e --no-cpu-baseline"
$CMDD > $O/plain_duke.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_duke.csv $CMDD > $O/ncu3.log 2>&1; echo "ncu launches duke rc=$?"
ls -la $O | head -20
