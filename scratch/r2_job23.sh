#!/bin/bash
set -u
O=gpurun_out/r2w; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-200
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > $O/bench_default.log 2>&1; echo "bench rc=$?"; grep "^{" $O/bench_default.log | cut -c1-400
timeout 600 python bench.py --impl reference > $O/bench_reference.log 2>&1; echo "ref rc=$?"; grep "^{" $O/bench_reference.log | cut -c1-400
