CMD2="python bench.py --workload msmt17 --steps 1 --warmup 1 --settle 0 --no-cpu-baseline --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'score_sparse' -s 1 -c 1 -o gpurun_out/prof_score $CMD2 > gpurun_out/ncu_score.log 2>&1
tail -n 2 gpurun_out/ncu_score.log
