#!/bin/bash
set -u
O=gpurun_out/r2p; mkdir -p $O
for P in 2 4; do
  REID_MR_SAME_GPU=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 2956$P tests/multi_rank_check.py > $O/mr_same_$P.log 2>&1
  echo "same-gpu multi_rank_check P=$P rc=$?"; grep -v "^W\|OMP_NUM\|\*\*\*" $O/mr_same_$P.log | tail -9 | cut -c1-300
done
