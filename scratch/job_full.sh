set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_duke.log 2>&1; tail -1 gpurun_out/b_duke.log | cut -c1-200
timeout 600 python bench.py --workload msmt17 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_msmt.log 2>&1; tail -1 gpurun_out/b_msmt.log | cut -c1-200
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_duke2.csv python bench.py --steps 2 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu1b.log 2>&1
