#!/bin/bash
# round 2, job 20 (EIGHT GPUs): graph-replayed multi-rank steps -- check at 8 ranks, small shapes at 4 / 8 ranks
set -u
O=gpurun_out/r2t; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
REID_MR_BIG=1 timeout 900 $TR --nproc-per-node 8 --master-port 29591 tests/multi_rank_check.py > $O/mr_8.log 2>&1; echo "multi_rank_check@8 rc=$?"
grep -v "$F" $O/mr_8.log | tail -9 | cut -c1-220
port=29592
for cfg in "duke 8" "duke 4" "market 8" "market 4" "veri 8" "veri 4" "msmt17 8"; do
  set -- $cfg; port=$((port+1))
  timeout 600 $TR --nproc-per-node $2 --master-port $port bench.py --gpus $2 --workload $1 --steps 20 --warmup 5 > $O/bench_$1_$2.log 2>&1
  echo "bench $1@$2 rc=$?"
  grep "^{" $O/bench_$1_$2.log | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('   ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_step'],3), 'enqueue', round(j['host_enqueue_ms_per_step'],3), 'launches', j['gpu_launches'], 'mAP', j['mAP'])"
done
