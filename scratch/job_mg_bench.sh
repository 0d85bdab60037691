N=${1:-4}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/mg_duke_$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --workload msmt17 2>/dev/null | tail -1 > gpurun_out/mg_msmt_$N.log
python - <<PY
import json
for f in ("gpurun_out/mg_duke_$N.log","gpurun_out/mg_msmt_$N.log"):
    j=json.loads(open(f).read()); print(f, j["n_gpus"], round(j["ms_per_step"],3), round(j["e2e"]["ms_per_step"],3), j["mAP"])
PY
