#!/bin/bash
set -u
O=gpurun_out/r2j; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -x --durations=8 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -14 $O/pytest.log | cut -c1-200
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --stages --no-cpu-baseline > $O/bench_duke.log 2>&1; tail -11 $O/bench_duke.log | cut -c1-420
timeout 300 python bench.py --workload duke --k1 50 --k2 15 --steps 20 --warmup 3 --stages --no-cpu-baseline > $O/bench_duke_50_15.log 2>&1; tail -11 $O/bench_duke_50_15.log | cut -c1-300
timeout 300 python bench.py --workload market --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_market.log 2>&1; tail -1 $O/bench_market.log | cut -c1-300
timeout 300 python bench.py --workload veri --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_veri.log 2>&1; tail -1 $O/bench_veri.log | cut -c1-300
timeout 400 python bench.py --steps 10 --warmup 3 --stages > $O/bench_msmt17.log 2>&1; tail -12 $O/bench_msmt17.log | cut -c1-600
