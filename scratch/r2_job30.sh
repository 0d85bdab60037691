#!/bin/bash
set -u
O=gpurun_out/r2ac; mkdir -p $O
for P in 2 3; do
  REID_MR_SAME_GPU=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 2958$P tests/multi_rank_check.py > $O/mr_same_$P.log 2>&1; echo "mr P=$P rc=$?"; grep -i "streamed\|error\|assert" $O/mr_same_$P.log | head -6 | cut -c1-250
done
timeout 900 python -m pytest tests/test_gpu_retrieval.py -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log | cut -c1-300
