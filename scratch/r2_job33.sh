#!/bin/bash
set -u
O=gpurun_out/r2af; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-200
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > $O/bench_default.log 2>&1; echo "bench rc=$?"; grep "^{" $O/bench_default.log | cut -c1-300
timeout 600 python bench.py --impl reference > $O/bench_reference.log 2>&1; echo "ref rc=$?"; grep "^{" $O/bench_reference.log | cut -c1-200
CMD="python bench_retrieval.py --q 20000 --g 200000 --rerank --steps 1 --warmup 1"
timeout 200 $CMD > $O/plain_20k_200k.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file $O/launches_rerank_20k_200k.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
tail -1 $O/plain_20k_200k.log | cut -c1-700
