#!/bin/bash
set -u
mkdir -p gpurun_out/last
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py > gpurun_out/last/bench.log 2>&1; tail -1 gpurun_out/last/bench.log | cut -c1-400
timeout 120 python bench.py --impl reference --steps 1 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
