#!/bin/bash
set -u
O=gpurun_out/r2r; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
show() { grep "^{" $1 | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('   ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_step'],3), 'enqueue', round(j['host_enqueue_ms_per_step'],3), 'launches', j['gpu_launches'])"; }
REID_GRAPH_BELOW=0 timeout 600 $TR --master-port 29581 bench.py --gpus 2 --workload duke --steps 20 --warmup 3 > $O/duke_nograph.log 2>&1; echo "no graph rc=$?"; show $O/duke_nograph.log
timeout 600 $TR --master-port 29582 bench.py --gpus 2 --workload duke --steps 20 --warmup 3 > $O/duke_graph.log 2>&1; echo "graph rc=$?"; show $O/duke_graph.log
timeout 600 $TR --master-port 29583 bench.py --gpus 2 --workload duke --steps 20 --warmup 3 --settle 0 > $O/duke_graph_nosettle.log 2>&1; echo "graph, settle 0 rc=$?"; show $O/duke_graph_nosettle.log
