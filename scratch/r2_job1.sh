#!/bin/bash
# round 2, first GPU call: whole -m gpu suite (incl. the new full-shape oracle parity), smoke, msmt17 bench, reference arm
set -u
O=gpurun_out/r2a; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > $O/gpu.txt; nproc >> $O/gpu.txt; free -g | head -2 >> $O/gpu.txt
timeout 1500 python -m pytest tests -q -m gpu -s --durations=15 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest.log
grep "\[parity\]" $O/pytest.log > $O/parity.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --steps 10 --warmup 3 --stages > $O/bench_msmt17.log 2>&1; tail -12 $O/bench_msmt17.log | cut -c1-1500
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --stages --no-cpu-baseline > $O/bench_duke.log 2>&1; tail -1 $O/bench_duke.log | cut -c1-600
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.log 2>&1; tail -1 $O/bench_ref.log | cut -c1-300
