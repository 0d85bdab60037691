#!/bin/bash
set -u
O=gpurun_out/r2m; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > $O/bench_msmt17_2.log 2>&1; grep "^{" $O/bench_msmt17_2.log | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('msmt17@2', j['ms_per_step'], 'enqueue', j['host_enqueue_ms_per_step'])"
timeout 600 $TR --master-port 29543 bench.py --gpus 2 --workload duke --steps 20 --warmup 3 --no-e2e > $O/bench_duke_2.log 2>&1; grep "^{" $O/bench_duke_2.log | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('duke@2', j['ms_per_step'], 'enqueue', j['host_enqueue_ms_per_step'])"
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_duke_1.log 2>&1; grep "^{" $O/bench_duke_1.log | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('duke@1', j['ms_per_step'], 'enqueue', j['host_enqueue_ms_per_step'], j['distmat_only'])"
