#!/bin/bash
# round 2, job 9 (EIGHT GPUs): both multi-rank paths == one rank at 8 ranks, MSMT17 / Duke benches at 4 and 8 ranks,
# BASELINE config 5 (100k x 1M) re-ranked and plain at 8 ranks
set -u
O=gpurun_out/r2i; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
F='^W\|OMP_NUM\|\*\*\*\*'
REID_MR_BIG=1 timeout 900 $TR --nproc-per-node 8 --master-port 29531 tests/multi_rank_check.py > $O/mr_8.log 2>&1; echo "multi_rank_check@8 rc=$?"
grep -v "$F" $O/mr_8.log | tail -18 | cut -c1-200
timeout 600 $TR --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 10 --warmup 3 --stages > $O/bench_msmt17_8.log 2>&1; echo "bench msmt17@8 rc=$?"
grep -v "$F" $O/bench_msmt17_8.log | tail -18 | cut -c1-1200
timeout 600 $TR --nproc-per-node 4 --master-port 29533 bench.py --gpus 4 --steps 10 --warmup 3 --stages > $O/bench_msmt17_4.log 2>&1; echo "bench msmt17@4 rc=$?"
grep -v "$F" $O/bench_msmt17_4.log | tail -16 | cut -c1-300
timeout 600 $TR --nproc-per-node 8 --master-port 29534 bench.py --gpus 8 --workload duke --steps 20 --warmup 3 --stages > $O/bench_duke_8.log 2>&1; echo "bench duke@8 rc=$?"
grep -v "$F" $O/bench_duke_8.log | tail -16 | cut -c1-300
timeout 900 $TR --nproc-per-node 8 --master-port 29535 bench_retrieval.py --rerank --steps 2 --warmup 1 > $O/bench_rerank_100k_1M_8.log 2>&1; echo "retrieval rerank@8 rc=$?"
grep -v "$F" $O/bench_rerank_100k_1M_8.log | tail -2 | cut -c1-1600
timeout 600 $TR --nproc-per-node 8 --master-port 29536 bench_retrieval.py --score --steps 3 --warmup 1 > $O/bench_score_100k_1M_8.log 2>&1; echo "retrieval score@8 rc=$?"
grep -v "$F" $O/bench_score_100k_1M_8.log | tail -1 | cut -c1-900
