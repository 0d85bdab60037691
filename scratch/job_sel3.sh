SHAPES=duke timeout 600 python scratch/sel_scatter.py > gpurun_out/plain_sel.log 2>&1 &&
SHAPES=duke timeout 900 ncu --set full --clock-control none --import-source on -k regex:'select_push|select_merge|select_thr' -s 3 -c 3 -o gpurun_out/prof_sel python scratch/sel_scatter.py > gpurun_out/ncu_sel.log 2>&1
tail -n 2 gpurun_out/ncu_sel.log
