#!/bin/bash
set -u
O=gpurun_out/r2v; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_retrieval.py -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-200
timeout 600 python bench_retrieval.py --rerank --steps 2 --warmup 1 > $O/bench_rerank_100k_1M.log 2>&1; tail -1 $O/bench_rerank_100k_1M.log | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('ms', round(j['ms_per_step'],1), {k:round(v,1) for k,v in j['stages_ms'].items()}, j['score'])"
