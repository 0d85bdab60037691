set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_duke.log 2>&1; tail -1 gpurun_out/b_duke.log | cut -c1-200
timeout 600 python bench.py --workload msmt17 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_msmt.log 2>&1; tail -1 gpurun_out/b_msmt.log | cut -c1-200
