"""Headline benchmark: re-ranked q x g pair-evaluations per second (distmat + k-reciprocal re-rank +
CMC/mAP) on synthetic features of a named dataset shape (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W [--workload duke|market|veri|msmt17]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      # CPU arm: the oracle port on the host cores

One "step" = one complete re-ranked evaluation of the workload: raw features -> L2 normalise ->
N x N squared-L2 -> k-reciprocal re-ranking (k1, k2, lambda of the config) -> q x g ranking ->
CMC/mAP.  `value` = q*g*K / device time with the features already resident in HBM; `e2e` = the same
through the drop-in evaluator (evaluate.R1_mAP_eval) from PINNED HOST buffers, H2D of features/ids
and D2H of (cmc, mAP) inside the timed region.  Multi-GPU: probe/query rows are sharded over the
ranks (strong scaling of one evaluation), time = max over ranks.

Timing: CUDA events around every step on the launching stream; between steps a 512 MB buffer is
rewritten to flush the 126 MB L2 (outside the per-step events).  nvidia-smi clocks are sampled
during the timed region.
"""
from __future__ import annotations

import argparse
import contextlib
import importlib
import io
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "re-ranked qxg pair-evals/sec (distmat+rerank+mAP)"
UNIT = "pair-evals/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]),
                    bf16_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback")


class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.samples:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
# CPU arm / cpu_baseline: the oracle port on a bounded, down-scaled sample of the same workload
# ----------------------------------------------------------------------------------------------
def cpu_sample_shape(synth, shape, scale: int):
    return synth.scaled(shape, max(8, shape.q // scale), max(64, shape.g // scale))


def run_cpu_oracle(synth, shape, seed=0):
    """One re-ranked evaluation by the oracle port (numpy; BLAS threads for the contraction).
    Returns (seconds, pairs, mAP)."""
    import torch
    from oracle import reid_oracle as oracle          # checker / CPU baseline only
    qf, gf, qp, gp, qc, gc = synth.make(shape, seed)
    feats = torch.cat([qf, gf]).numpy()
    pids = torch.cat([qp, gp]).numpy()
    cams = torch.cat([qc, gc]).numpy()
    t0 = time.perf_counter()
    cmc, mAP = oracle.evaluate_features(feats, pids, cams, shape.q, feat_norm=True, reranking=True,
                                        k1=shape.k1, k2=shape.k2, lam=shape.lam)
    dt = time.perf_counter() - t0
    return dt, shape.q * shape.g, float(mAP)


def cpu_threads():
    try:
        import torch
        return int(torch.get_num_threads())
    except Exception:
        return os.cpu_count() or 1


def reference_arm(args, synth, shape, rank):
    if rank != 0:
        return
    sample = cpu_sample_shape(synth, shape, args.cpu_scale)
    for _ in range(max(0, min(args.warmup, 1))):
        run_cpu_oracle(synth, sample)
    times, pairs = [], 0
    for _ in range(args.steps):
        dt, pairs, mAP = run_cpu_oracle(synth, sample)
        times.append(dt)
    total = sum(times)
    value = pairs * len(times) / total
    desc = (f"{sample.q}x{sample.g}x{sample.d} (1/{args.cpu_scale} of {shape.name} in q and g, same generator), "
            f"k1={sample.k1} k2={sample.k2} lambda={sample.lam}; full re-ranked eval per step")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": shape.name, "q": shape.q, "g": shape.g, "d": shape.d, "k1": shape.k1, "k2": shape.k2,
                   "lambda": shape.lam, "sample": desc},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": os.cpu_count(), "threads_blas": cpu_threads(),
                         "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="duke")
    ap.add_argument("--engine", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--cpu-scale", type=int, default=2, help="CPU sample = workload with q and g divided by this")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--stages", action="store_true", help="print a per-stage table to stderr")
    ap.add_argument("--settle", type=int, default=100,
                    help="extra untimed steps before the timed region so that nvidia-smi's 100 ms sampling sees load")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    synth = importlib.import_module("multimodal-reid_b200.synth")
    shape = synth.SHAPES[args.workload]

    if args.impl == "reference":
        reference_arm(args, synth, shape, rank)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl b200) needs a B200; there is no CPU fallback. Use --impl reference for the CPU arm.")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    engine = importlib.import_module("multimodal-reid_b200.engine")
    evaluate = importlib.import_module("multimodal-reid_b200.evaluate")
    _lib = importlib.import_module("multimodal-reid_b200._lib")
    eng = engine.Engine(dev, engine=args.engine)
    engine._default = eng
    lib = _lib.load()
    peaks = load_peaks()

    # ---- synthetic workload, identical on every rank
    qf, gf, qp, gp, qc, gc = synth.make(shape, 0)
    q, g, d = shape.q, shape.g, shape.d
    N = q + g
    feats_h = torch.cat([qf, gf]).pin_memory()
    pids_h = torch.cat([qp, gp]).pin_memory()
    cams_h = torch.cat([qc, gc]).pin_memory()
    feats_d = feats_h.to(dev)
    pids_d, cams_d = pids_h.to(dev), cams_h.to(dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    cm = eng.comm
    qsel = slice(cm.rank, q, cm.world)

    def step_resident():
        cmc, mAP, nvalid, _, _ = eng.evaluate_reranked(feats_d, q, pids_d, cams_d, shape.k1, shape.k2, shape.lam,
                                                       normalize=True, max_rank=50)
        return cmc, mAP

    def step_e2e():
        e = evaluate.R1_mAP_eval(q, max_rank=50, feat_norm=True, reranking=True,
                                 rerank_params=(shape.k1, shape.k2, shape.lam))
        e.reset()
        e.update((feats_h, pids_h, cams_h))
        with contextlib.redirect_stdout(io.StringIO()):
            return e.compute()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, timers=False):
        evs = []
        barrier()
        t_wall = time.perf_counter()
        for _ in range(steps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if timers:
                eng.enable_timers(True)
            a.record()
            out = fn()
            b.record()
            evs.append((a, b))
        barrier()
        wall = time.perf_counter() - t_wall
        ms = [a.elapsed_time(b) for a, b in evs]
        tot = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        return float(tot.item()), ms, out, wall

    # ---- warm-up (the clock sampler starts here: nvidia-smi needs ~100 ms to deliver its first line, and the
    #      timed region of this workload is only tens of milliseconds; warm-up runs the same kernels)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    n_warm = max(args.warmup, 3) + args.settle       # same count on every rank (the steps contain collectives)
    for _ in range(n_warm):
        out = step_resident()
    barrier()

    # ---- timed: device-resident
    l0 = lib.reid_launch_count()
    total_ms, per_step, out, wall = timed(step_resident, args.steps)
    launches = int(lib.reid_launch_count() - l0)
    cmc, mAP = out
    ms_per_step = total_ms / args.steps
    value = q * g * args.steps / (total_ms * 1e-3)

    # ---- per-stage timing of the same step (separate pass; events on the launching stream)
    eng.enable_timers(True)
    stage_steps = min(args.steps, 5)
    barrier()
    for _ in range(stage_steps):
        flush.fill_(1)
        step_resident()
    stages = {k: v / stage_steps for k, v in eng.stage_ms().items()}
    eng.enable_timers(False)

    # ---- e2e through the drop-in evaluator, host buffers
    e2e = None
    if not args.no_e2e:
        for _ in range(2):
            step_e2e()
        e2e_ms, _, (cmc_e, map_e), _ = timed(step_e2e, args.steps)
        # several ranks: each uploads its 1/world block of the features (Comm.upload_sharded) and all the ids
        ids_b = pids_h.numel() * 8 + cams_h.numel() * 8
        lo_r, hi_r = cm.block_bounds(N)
        h2d_rank = (hi_r - lo_r) * d * 4 + ids_b if world > 1 else feats_h.numel() * 4 + ids_b
        h2d = feats_h.numel() * 4 + world * ids_b             # whole job
        d2h = 50 * 4 + 8 + 4 + 8          # cmc, mAP, n_valid, longest-union hint (async, pinned)
        e2e = {"value": q * g * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
               "h2d_bytes_per_step": h2d, "h2d_bytes_per_step_this_rank": h2d_rank, "d2h_bytes_per_step": d2h,
               "mAP": float(map_e),
               "api": "evaluate.R1_mAP_eval(reranking=True).update(host tensors).compute()"}

    clocks = sampler.stop() if rank == 0 else None

    # ---- roofline of the dominant kernel (N x N distance contraction)
    rows_local = cm.local_count(N)
    flops = 2.0 * rows_local * N * d          # algorithmic, per launch (this rank's panel): SURVEY.md 8(d)
    dist_ms = stages.get("dist", float("nan"))
    # burst figure when the clocks stayed at their maximum over the timed region (short steps: Duke, Market, VeRi), the
    # sustained one when the chip lowered them under the load (MSMT17: ~1.44 GHz, sw_power_cap)
    at_max = bool(clocks and clocks.get("sm_mhz") and clocks.get("sm_max_mhz")
                  and clocks["sm_mhz"] >= 0.97 * clocks["sm_max_mhz"])
    which = "bf16" if at_max else "bf16_sustained"
    bf16_peak = peaks[which]
    tf32_peak = bf16_peak / 2.0
    achieved = flops / (dist_ms * 1e-3) / 1e12
    # what the tensor pipe is actually asked to do: 3 fp16 MMAs (hi.hi + hi.lo + lo.hi) over the 256 x 256 tiles the
    # schedule computes -- the upper triangle only when one rank owns every row (symmetric mode)
    T_n, dp = -(-N // 256), -(-d // 64) * 64
    tiles = T_n * (T_n + 1) // 2 if (world == 1 and eng.symmetric) else -(-rows_local // 256) * T_n
    issued = 3.0 * tiles * 2.0 * 256 * 256 * dp
    issued_tf = issued / (dist_ms * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if world == 1 and os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(shape.name, {}).get("dist_tc2_kernel")
    roofline = {"kernel": "dist_tc2_kernel (N x N squared-L2 contraction, tcgen05 cta_group::2)", "bound": "tensor",
                "achieved": achieved, "peak": tf32_peak, "unit": "TFLOP/s", "frac": achieved / tf32_peak,
                "peak_source": f"{peaks['source']}: {which}/2 = TF32-equivalent dense ("
                               + ("burst: SM clocks stayed at max" if at_max else "sustained: SM clocks lowered under load")
                               + ")",
                "frac_note": "achieved counts the 2*rows*N*d flops of the full matrix (SURVEY 8d); with one rank the "
                             "symmetric schedule computes only the upper-triangle tiles and mirrors the rest, so frac can "
                             "exceed 1 -- the tensor-pipe fraction is issued_frac_of_bf16_peak",
                "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu --set full, dram read+write per launch)",
                "ms": dist_ms, "algorithmic_flops_per_launch": flops,
                "issued_fp16_tflops": issued_tf, "issued_frac_of_bf16_peak": issued_tf / bf16_peak,
                "tiles_256x256": tiles, "symmetric_schedule": bool(world == 1 and eng.symmetric)}
    hbm = {}
    # algorithmic bytes: selection reads the chunk minima (+ ~K candidate lines per row); the fused
    # Jaccard/score kernel streams its queries' rows of D once and writes nothing q x g
    K = max(shape.k1 + 1, shape.k2)
    bytes_stage = {"select": 4.0 * rows_local * ((N + 31) // 32) + 128.0 * K * rows_local,
                   "jaccard_score": 4.0 * cm.local_count(q) * N}
    for k, b in bytes_stage.items():
        if k in stages and stages[k] > 0:
            gbs = b / (stages[k] * 1e-3) / 1e9
            hbm[k] = {"GB/s": gbs, "frac": gbs / peaks["hbm"]}

    # ---- CPU baseline (rank 0, N == 1 only): oracle port on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = cpu_sample_shape(synth, shape, args.cpu_scale)
        dt, pairs, m = run_cpu_oracle(synth, sample)
        cpu = {"value": pairs / dt, "unit": UNIT, "cores": os.cpu_count(), "threads_blas": cpu_threads(),
               "kind": "port", "seconds": dt,
               "sample": f"{sample.q}x{sample.g}x{sample.d} (1/{args.cpu_scale} of {shape.name} in q and g), "
                         f"k1={sample.k1} k2={sample.k2}; one full re-ranked eval by oracle/reid_oracle.py"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": n_warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": shape.name, "q": q, "g": g, "d": d, "k1": shape.k1, "k2": shape.k2,
                       "lambda": shape.lam, "sharding": f"probe rows cyclic over {world} rank(s)",
                       "l2_flush_between_steps": True, "engine": args.engine},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "stages_ms": stages, "hbm_stages": hbm, "mAP": float(mAP.item()), "rank1": float(cmc[0].item()),
            "wall_s_timed_region": wall,
        }
        print(json.dumps(line), flush=True)
        if args.stages:
            for k, v in sorted(stages.items(), key=lambda kv: -kv[1]):
                print(f"  {k:14s} {v:9.3f} ms", file=sys.stderr)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
