#!/bin/bash
set -u
O=gpurun_out/r2s; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-200
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_duke.log 2>&1; grep "^{" $O/bench_duke.log | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('duke@1 ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_step'],3), 'enqueue', round(j['host_enqueue_ms_per_step'],3), 'launches', j['gpu_launches'])"
