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
# CPU arm / cpu_baseline: the reference's own CPU implementation of the path on the box's host cores.
#   kind "reference": the UNMODIFIED /root/reference/{evaluate,reranking}.py, shipped by oracle/build_ref.py as
#                     oracle/_ref (git-ignored, travels with the snapshot): re_ranking(...) + eval_func(...), the body of
#                     R1_mAP_eval(reranking=True).compute() (evaluate.py:110-134) with the config's k1/k2/lambda
#   kind "port":      oracle/reid_oracle.py (sparse numpy restatement, several times FASTER than the reference) -- only
#                     when oracle/_ref did not arrive
# The reference keeps ~24 N^2 bytes live and argsorts the full N x N matrix (reranking.py:48): 66 s at N = 19 k, ~210 GB
# at the MSMT17 shape.  A step of the CPU arm is therefore a bounded, down-scaled sample of the workload (same generator,
# same q:g ratio and identity density); the shape that ran is named in `sample`.
# ----------------------------------------------------------------------------------------------
CPU_SCALE_DEFAULT = {"msmt17": 16, "duke": 4, "market": 4, "veri": 2}


def cpu_sample_shape(synth, shape, scale: int):
    if scale <= 1:
        return shape
    return synth.scaled(shape, max(8, shape.q // scale), max(64, shape.g // scale))


def cpu_scale_for(args, shape) -> int:
    return args.cpu_scale if args.cpu_scale > 0 else CPU_SCALE_DEFAULT.get(shape.name, 4)


def cpu_kind() -> str:
    from oracle import build_ref                      # checker / CPU baseline only
    return "reference" if build_ref.available() else "port"


def run_cpu_eval(synth, shape, seed=0):
    """One re-ranked evaluation on the host cores.  Returns (seconds, pairs, mAP, kind)."""
    import torch
    qf, gf, qp, gp, qc, gc = synth.make(shape, seed)
    kind = cpu_kind()
    if kind == "reference":
        from oracle import build_ref
        ev, rr = build_ref.load()
        t0 = time.perf_counter()
        # evaluate.py:110-134 with reranking=True (k1/k2/lambda of the config instead of the hard-coded 50/15/0.3)
        feats = torch.nn.functional.normalize(torch.cat([qf, gf]), dim=1, p=2)
        with contextlib.redirect_stdout(io.StringIO()):
            distmat = rr.re_ranking(feats[: shape.q], feats[shape.q:], k1=shape.k1, k2=shape.k2, lambda_value=shape.lam)
            cmc, mAP = ev.eval_func(distmat, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
        dt = time.perf_counter() - t0
    else:
        from oracle import reid_oracle as oracle
        feats = torch.cat([qf, gf]).numpy()
        pids = torch.cat([qp, gp]).numpy()
        cams = torch.cat([qc, gc]).numpy()
        t0 = time.perf_counter()
        cmc, mAP = oracle.evaluate_features(feats, pids, cams, shape.q, feat_norm=True, reranking=True,
                                            k1=shape.k1, k2=shape.k2, lam=shape.lam)
        dt = time.perf_counter() - t0
    return dt, shape.q * shape.g, float(mAP), kind


def cpu_threads():
    try:
        import torch
        return int(torch.get_num_threads())
    except Exception:
        return os.cpu_count() or 1


def cpu_sample_desc(sample, shape, scale, kind):
    what = ("unmodified reference re_ranking + eval_func (oracle/_ref)" if kind == "reference"
            else "oracle/reid_oracle.py (sparse numpy port; oracle/_ref absent)")
    if scale <= 1:
        return (f"{sample.q}x{sample.g}x{sample.d} = the full {shape.name} shape, k1={sample.k1} k2={sample.k2} "
                f"lambda={sample.lam}; one full re-ranked eval per step by the {what}")
    N = shape.q + shape.g
    return (f"{sample.q}x{sample.g}x{sample.d} (1/{scale} of {shape.name} in q and g, same generator), k1={sample.k1} "
            f"k2={sample.k2} lambda={sample.lam}; one full re-ranked eval per step by the {what}; the full shape needs "
            f"~24*N^2 B = {24 * N * N / 1e9:.0f} GB and a full N x N argsort per step on the host")


def reference_arm(args, synth, shape, rank):
    if rank != 0:
        return
    scale = cpu_scale_for(args, shape)
    sample = cpu_sample_shape(synth, shape, scale)
    for _ in range(max(0, min(args.warmup, 1))):
        run_cpu_eval(synth, sample)
    times, pairs, kind = [], 0, "port"
    for _ in range(args.steps):
        dt, pairs, mAP, kind = run_cpu_eval(synth, sample)
        times.append(dt)
    total = sum(times)
    value = pairs * len(times) / total
    desc = cpu_sample_desc(sample, shape, scale, kind)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": shape.name, "q": shape.q, "g": shape.g, "d": shape.d, "k1": shape.k1, "k2": shape.k2,
                   "lambda": shape.lam, "sample": desc, "sample_q": sample.q, "sample_g": sample.g,
                   "same_config": scale <= 1},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": os.cpu_count(), "threads_blas": cpu_threads(),
                         "kind": kind, "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "mAP_sample": mAP,
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
    ap.add_argument("--workload", default="msmt17", help="msmt17 (the north-star shape, default) | duke | market | veri")
    ap.add_argument("--k1", type=int, default=0, help="override the workload's k1 (e.g. 50, the evaluator's own value)")
    ap.add_argument("--k2", type=int, default=0)
    ap.add_argument("--engine", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--cpu-scale", type=int, default=0,
                    help="CPU sample = workload with q and g divided by this (0: per-workload default; 1: full shape)")
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
    if args.k1 or args.k2:
        import dataclasses
        shape = dataclasses.replace(shape, k1=args.k1 or shape.k1, k2=args.k2 or shape.k2)

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
                                                       normalize=True, max_rank=50, resident=True)
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

    host_enqueue = []       # host time spent issuing one step (no synchronisation inside): launch-bound when close to the step time

    def timed(fn, steps, timers=False):
        evs = []
        host_enqueue.clear()
        barrier()
        t_wall = time.perf_counter()
        for _ in range(steps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if timers:
                eng.enable_timers(True)
            a.record()
            t_h = time.perf_counter()
            out = fn()
            host_enqueue.append(time.perf_counter() - t_h)
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
    n_warm = max(args.warmup, 5) + args.settle       # same count on every rank; >= 5 so that a graph capture (4th call) is done
    for _ in range(n_warm):
        out = step_resident()
    barrier()

    # ---- timed: device-resident
    l0 = lib.reid_launch_count()
    total_ms, per_step, out, wall = timed(step_resident, args.steps)
    launches = int(lib.reid_launch_count() - l0)
    enqueue_ms = 1e3 * statistics.median(host_enqueue) if host_enqueue else None
    cmc, mAP = out
    ms_per_step = total_ms / args.steps
    value = q * g * args.steps / (total_ms * 1e-3)

    # ---- per-stage timing of the same step (separate pass; events on the launching stream)
    eng.enable_timers(True)
    stage_steps = min(args.steps, 5)
    for _ in range(3):                 # with timers on, one rank takes the staged path (not the single C entry): warm its
        step_resident()                # allocations before timing the stages
    eng.enable_timers(True)            # (drops the warm-up spans)
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

    # ---- BASELINE configs[1] names both distance functions at the Duke shape: the q x g matrices of evaluate.py:7-13 and
    #      :16-26 and eval_func on them (device-resident, CUDA events, one rank), reported beside the headline
    distmat = None
    if world == 1 and N <= 40_000:
        allf = eng.prepare(feats_d, True)
        Qf, Gf = allf.rows(slice(0, q)), allf.rows(slice(q, N))
        idq = (pids_d[:q].contiguous(), pids_d[q:].contiguous(), cams_d[:q].contiguous(), cams_d[q:].contiguous())

        def timed_op(fn, reps=10):
            for _ in range(3):
                fn()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush.fill_(1)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize(dev)
            return a.elapsed_time(b) / reps

        De, _ = eng.dist(Qf, Gf, engine.METRIC_SQEUCLIDEAN)
        distmat = {"euclidean_distance_ms": timed_op(lambda: eng.dist(Qf, Gf, engine.METRIC_SQEUCLIDEAN)),
                   "cosine_similarity_ms": timed_op(lambda: eng.dist(Qf, Gf, engine.METRIC_ARCCOS)),
                   "eval_func_ms": timed_op(lambda: eng.evaluate_dist(De, g, *idq, 50)),
                   "note": "q x g matrix on the device (evaluate.py:7-13 / :16-26) and CMC/mAP of it (evaluate.py:29-88); "
                           "no re-ranking, no host copies"}
        del De, allf

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
    fabric_path = world > 1 and eng.last_path == "fabric"
    sym_sched = eng.symmetric and (world == 1 or fabric_path)
    tiles = eng.last_tiles if fabric_path else (T_n * (T_n + 1) // 2 if sym_sched else -(-rows_local // 256) * T_n)
    if fabric_path:
        flops = 2.0 * N * N * d / world       # this rank's share of the symmetric matrix (snake tile rows)
        achieved = flops / (dist_ms * 1e-3) / 1e12
    issued = 3.0 * tiles * 2.0 * 256 * 256 * dp
    issued_tf = issued / (dist_ms * 1e-3) / 1e12
    traffic_all = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic_all = json.load(f).get(shape.name, {})
    traffic = traffic_all.get("dist_tc2_kernel") if world == 1 else None
    # frac = what the tensor pipe executes / what it can execute: issued fp16 MMA flops (3 per 256 x 256 x dp tile the
    # schedule really computes) over the measured dense bf16 rate (kind::f16 MMAs run at the bf16 rate).  The
    # algorithmic 2*rows*N*d figure of SURVEY 8(d) is reported beside it, against the TF32-equivalent (bf16 / 2) peak; it
    # can exceed that peak because the symmetric schedule computes only the upper-triangle tiles.
    roofline = {"kernel": "dist_tc2_kernel (N x N squared-L2 contraction, tcgen05 cta_group::2, fp16 hi/lo split)",
                "bound": "tensor", "achieved": issued_tf, "peak": bf16_peak, "unit": "TFLOP/s",
                "frac": issued_tf / bf16_peak,
                "peak_source": f"{peaks['source']}: {which} ("
                               + ("burst: SM clocks stayed at max" if at_max else "sustained: SM clocks lowered under load")
                               + ")",
                "achieved_is": "issued fp16 MMA flops = 3 * tiles * 2*256*256*dp per launch / live CUDA-event time",
                "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu --set full, dram read+write per launch)",
                "ms": dist_ms, "tiles_256x256": tiles, "symmetric_schedule": bool(sym_sched),
                "algorithmic_flops_per_launch": flops, "algorithmic_tflops": achieved,
                "algorithmic_frac_of_tf32_equiv_peak": achieved / tf32_peak}
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

    # ncu-measured DRAM bytes of the HBM-bound kernels (profiles/traffic.json, one launch each) with their ncu durations are
    # in profiles/r02k_kernels_msmt17_key_metrics.txt; here: the stage-level algorithmic figures above plus those bytes
    for kname in ("score_sparse_kernel", "select_push_kernel", "overlap_sparse_kernel"):
        if world == 1 and kname in traffic_all:
            hbm.setdefault("ncu_dram_bytes_per_launch", {})[kname] = traffic_all[kname]

    # ---- CPU baseline (rank 0, N == 1 only): oracle port on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        scale = cpu_scale_for(args, shape)
        sample = cpu_sample_shape(synth, shape, scale)
        dt, pairs, m, kind = run_cpu_eval(synth, sample)
        cpu = {"value": pairs / dt, "unit": UNIT, "cores": os.cpu_count(), "threads_blas": cpu_threads(),
               "kind": kind, "seconds": dt, "sample": cpu_sample_desc(sample, shape, scale, kind)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": n_warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": shape.name, "q": q, "g": g, "d": d, "k1": shape.k1, "k2": shape.k2,
                       "lambda": shape.lam,
                       "sharding": ("one rank" if world == 1 else
                                    f"symmetric contraction sharded by snake tile rows over {world} ranks, exchanges through "
                                    "peer memory (CUDA IPC over NVLink), flag barriers" if fabric_path else
                                    f"probe rows cyclic over {world} ranks, NCCL all-gathers"),
                       "l2_flush_between_steps": True, "engine": args.engine},
            "e2e": e2e, "gpu_launches": launches, "host_enqueue_ms_per_step": enqueue_ms, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "stages_ms": stages, "hbm_stages": hbm, "distmat_only": distmat, "mAP": float(mAP.item()),
            "rank1": float(cmc[0].item()),
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
