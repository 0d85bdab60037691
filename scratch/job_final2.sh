#!/bin/bash
set -u
mkdir -p gpurun_out/fin
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python bench_losses.py > gpurun_out/fin/bench_losses.log 2>&1; tail -6 gpurun_out/fin/bench_losses.log | cut -c1-220
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/fin/b1.log; python -c "
import json; j=json.loads(open('gpurun_out/fin/b1.log').read()); print(j['ms_per_step'], j['e2e']['ms_per_step'], {k:j['roofline'][k] for k in ('achieved','peak','frac','peak_source','issued_frac_of_bf16_peak','traffic')})"
bash scratch/job_mg.sh 2
