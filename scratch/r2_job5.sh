#!/bin/bash
# round 2, job 5 (ONE GPU): peer-memory path with 8 ranks time-slicing cuda:0 (functional: P = 8 index arithmetic,
# 8 x K merge, empty ranks at tiny shapes), incl. the Duke shape; then the whole -m gpu suite after the score-kernel revert
set -u
O=gpurun_out/r2e; mkdir -p $O
REID_MR_SAME_GPU=1 REID_MR_BIG=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 tests/multi_rank_check.py > $O/mr_same_8.log 2>&1
echo "same-gpu multi_rank_check P=8 rc=$?"; grep -v "^W\|OMP_NUM\|\*\*\*" $O/mr_same_8.log | tail -14 | cut -c1-300
timeout 1500 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
