#!/bin/bash
set -u
O=gpurun_out/r2u; mkdir -p $O
CMD="python bench_retrieval.py --q 20000 --g 200000 --rerank --steps 1 --warmup 1"
$CMD > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file $O/launches_rerank_20k_200k.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
tail -1 $O/plain.log | cut -c1-900
timeout 600 python bench_retrieval.py --rerank --steps 2 --warmup 1 > $O/bench_rerank_100k_1M.log 2>&1; tail -1 $O/bench_rerank_100k_1M.log | cut -c1-1200
