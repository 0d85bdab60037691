#!/bin/bash
set -u
O=gpurun_out/r2aj; mkdir -p $O
CMD="python bench_retrieval.py --q 20000 --g 200000 --rerank --steps 1 --warmup 1"
timeout 60 $CMD > $O/plain.log 2>&1 && timeout 100 ncu --set full --clock-control none -k regex:"overlap_hash_kernel|count_panel_rr_kernel" -c 3 --csv --page raw --log-file $O/a13_kernels_raw.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
ls -la $O | tail -3
