#!/bin/bash
set -u
O=gpurun_out/r2ai; mkdir -p $O
CMD="python bench_retrieval.py --q 20000 --g 200000 --rerank --steps 1 --warmup 1"
timeout 60 $CMD > $O/plain.log 2>&1 && timeout 100 ncu --set full --clock-control none -k regex:"sym_push_kernel|sym_fold_kernel" -s 20 -c 4 --csv --page raw --log-file $O/sym_kernels_raw.csv $CMD > $O/ncu.log 2>&1; echo "ncu rc=$?"
ls -la $O | tail -4
