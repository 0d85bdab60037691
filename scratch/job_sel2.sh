timeout 600 python -m pytest tests/test_gpu_stages.py -x -q -m gpu -k "select" 2>&1 | tail -2
for pf in 1 2; do echo pf=$pf; REID_SEL_PF=$pf timeout 600 python scratch/sel_scatter.py 2>&1 | tail -2; done
SHAPES=duke,msmt17 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum --clock-control none -k regex:select_ --csv --log-file gpurun_out/sel_launches.csv python scratch/sel_scatter.py > gpurun_out/sel_ncu.log 2>&1
