set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_duke.log 2>&1; tail -1 gpurun_out/b_duke.log | cut -c1-200
timeout 600 python bench.py --workload msmt17 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_msmt.log 2>&1; tail -1 gpurun_out/b_msmt.log | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum --clock-control none -k regex:'score_sparse|overlap_sparse' --csv --log-file gpurun_out/js_launches.csv python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e > gpurun_out/js_ncu.log 2>&1
