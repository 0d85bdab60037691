#!/bin/bash
set -u
O=gpurun_out/r2ah; mkdir -p $O
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 150 python bench.py > $O/bench_default.log 2>&1; echo "bench rc=$?"; grep "^{" $O/bench_default.log | cut -c1-260
