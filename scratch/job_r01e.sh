#!/bin/bash
# r01e: tests + bench (both arms) + ncu evidence of the final build, one B200
set -u
mkdir -p gpurun_out/r01e
O=gpurun_out/r01e
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py --stages > $O/b_duke.log 2>&1; tail -1 $O/b_duke.log | cut -c1-600
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/b_ref.log 2>&1; tail -1 $O/b_ref.log | cut -c1-400
timeout 600 python bench.py --workload msmt17 --steps 5 --warmup 3 --no-cpu-baseline --stages > $O/b_msmt.log 2>&1; tail -1 $O/b_msmt.log | cut -c1-400
for w in market veri; do timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $O/b_$w.log 2>&1; tail -1 $O/b_$w.log | cut -c1-200; done
CMD="python bench.py --steps 2 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
$CMD > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/launches_duke.csv $CMD > $O/ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dist_tc2_kernel -s 2 -c 1 -o $O/prof_tc2 $CMD > $O/ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'score_sparse|overlap_sparse|select_|qe_expand|expand_weights' -s 8 -c 8 -o $O/prof_duke_rr $CMD > $O/ncu2b.log 2>&1
$CMD2 > $O/plain_msmt.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_msmt.csv $CMD2 > $O/ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'dist_tc2_kernel|score_sparse|overlap_sparse|select_' -s 6 -c 6 -o $O/prof_msmt $CMD2 > $O/ncu4.log 2>&1
tail -n 1 $O/ncu1.log $O/ncu2.log $O/ncu2b.log $O/ncu3.log $O/ncu4.log
ls -la $O
