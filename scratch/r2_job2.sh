#!/bin/bash
# round 2, job 2 (ONE GPU): refactored kernels under the whole -m gpu suite, then the peer-memory path with 2 / 4 ranks
# time-slicing cuda:0 (REID_MR_SAME_GPU=1: gloo rendezvous, CUDA IPC arenas, flag barriers)
set -u
O=gpurun_out/r2b; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu -s --durations=10 -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.log
grep "\[parity\]" $O/pytest.log > $O/parity.txt; cat $O/parity.txt | cut -c1-260
for P in 2 4; do
  REID_MR_SAME_GPU=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 2951$P tests/multi_rank_check.py > $O/mr_same_$P.log 2>&1
  echo "same-gpu multi_rank_check P=$P rc=$?"; grep -v "^W\|OMP_NUM\|\*\*\*" $O/mr_same_$P.log | tail -14 | cut -c1-300
done
