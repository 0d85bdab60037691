#!/bin/bash
set -u
O=gpurun_out/r2n; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_stages.py tests/test_gpu_api.py tests/test_gpu_fullsize.py tests/test_gpu_retrieval.py -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-200
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 10 --warmup 3 --stages --no-cpu-baseline --no-e2e 2>&1 | tail -10 | cut -c1-200
