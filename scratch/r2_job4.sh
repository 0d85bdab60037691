#!/bin/bash
# round 2, job 4 (ONE GPU): score_sparse rewrite -- stage tests, full-size properties, benches with stage table, launch list
set -u
O=gpurun_out/r2d; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_stages.py tests/test_gpu_api.py tests/test_gpu_fullsize.py -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
timeout 400 python bench.py --steps 10 --warmup 3 --stages --no-cpu-baseline > $O/bench_msmt17.log 2>&1; tail -14 $O/bench_msmt17.log | cut -c1-700
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --stages --no-cpu-baseline > $O/bench_duke.log 2>&1; tail -12 $O/bench_duke.log | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -c 300 --csv --log-file $O/launches_msmt17.csv python bench.py --steps 2 --warmup 3 --settle 0 --no-e2e --no-cpu-baseline > $O/ncu.log 2>&1; echo "ncu rc=$?"
