#!/bin/bash
# round 2, job 15 (EIGHT GPUs): final build -- every named shape at 2 / 4 / 8 ranks, config 5 at 8 ranks
set -u
O=gpurun_out/r2o; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
port=29550
for cfg in "msmt17 8 10" "msmt17 4 10" "msmt17 2 10" "duke 8 20" "duke 4 20" "market 8 20" "market 4 20" "market 2 20" "veri 8 20" "veri 4 20" "veri 2 20"; do
  set -- $cfg; port=$((port+1))
  timeout 600 $TR --nproc-per-node $2 --master-port $port bench.py --gpus $2 --workload $1 --steps $3 --warmup 3 --stages > $O/bench_$1_$2.log 2>&1
  echo "bench $1@$2 rc=$?"
  grep "^{" $O/bench_$1_$2.log | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('   ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_step'],3), 'enqueue', round(j['host_enqueue_ms_per_step'],3), 'mAP', j['mAP'], {k:round(v,3) for k,v in j['stages_ms'].items()})"
done
timeout 900 $TR --nproc-per-node 8 --master-port 29570 bench_retrieval.py --rerank --steps 2 --warmup 1 > $O/bench_rerank_100k_1M_8.log 2>&1; echo "retrieval rerank@8 rc=$?"
grep "^{" $O/bench_rerank_100k_1M_8.log | cut -c1-300
