#!/bin/bash
set -u
O=gpurun_out/r2q; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
timeout 900 $TR --master-port 29571 tests/multi_rank_check.py > $O/mr_2.log 2>&1; echo "multi_rank_check rc=$?"; grep -v "$F" $O/mr_2.log | tail -8 | cut -c1-260
port=29572
for cfg in "duke 20" "market 20" "veri 20" "msmt17 10"; do
  set -- $cfg; port=$((port+1))
  timeout 600 $TR --master-port $port bench.py --gpus 2 --workload $1 --steps $2 --warmup 3 > $O/bench_$1_2.log 2>&1; echo "bench $1@2 rc=$?"
  grep "^{" $O/bench_$1_2.log | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('   ms', round(j['ms_per_step'],3), 'e2e', round(j['e2e']['ms_per_step'],3), 'enqueue', round(j['host_enqueue_ms_per_step'],3), 'launches', j['gpu_launches'], 'mAP', j['mAP'])"
done
