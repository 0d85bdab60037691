#!/bin/bash
# round 2, call 2 (ONE GPU): refactor sanity + the peer-memory path functionally, every rank on cuda:0 (gloo rendezvous)
set -u
O=gpurun_out/r2b; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_stages.py tests/test_gpu_api.py -x -q > $O/pytest_quick.log 2>&1; echo "quick rc=$?"; tail -3 $O/pytest_quick.log
for P in 2 4; do
  REID_MR_SAME_GPU=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 2951$P tests/multi_rank_check.py > $O/mr_same_$P.log 2>&1
  echo "same-gpu P=$P rc=$?"; grep -v "^\*\*\*\|OMP_NUM\|^$\|W1\|Setting" $O/mr_same_$P.log | tail -12
done
timeout 1500 python -m pytest tests -q -m gpu -s --durations=10 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest.log
grep "\[parity\]" $O/pytest.log > $O/parity.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_msmt17.log 2>&1; tail -1 $O/bench_msmt17.log | cut -c1-300
