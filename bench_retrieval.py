"""Streamed top-k retrieval at BASELINE.json's config 5 scale (first slice): 100k query x 1M gallery x
512-d, top-100, the q x g matrix (400 GB) never materialised.

    python bench_retrieval.py [--q 100000 --g 1000000 --d 512 --topk 100 --steps 3 --warmup 1]
    python -m torch.distributed.run --nproc-per-node N ... bench_retrieval.py      # queries sharded cyclically

Synthetic features are drawn on the device (seeded), L2-normalised by the same prepare kernel the evaluator uses.
Timed region (CUDA events, max over ranks): prepared features resident -> top-k lists resident (+ the all-gather of the
lists when N > 1).  One JSON line; `roofline` is the contraction's algorithmic 2*q*g*d flops against the TF32-equivalent
peak, `hbm_panel` the panel bytes the design still writes and re-reads (what a fused top-k epilogue would remove)."""
import argparse
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--q", type=int, default=100_000)
    ap.add_argument("--g", type=int, default=1_000_000)
    ap.add_argument("--d", type=int, default=512)
    ap.add_argument("--topk", type=int, default=100)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--rows-per-pass", type=int, default=16384)
    ap.add_argument("--cols-per-pass", type=int, default=65536)
    ap.add_argument("--score", action="store_true", help="also CMC / mAP over the streamed problem (same sweep; match distances contracted as gathered pairs first)")
    ap.add_argument("--rerank", action="store_true",
                    help="k-reciprocal re-ranked CMC / mAP + exact top-k of the re-ranked rows (N x N sweep + one q x g sweep)")
    ap.add_argument("--k1", type=int, default=20)
    ap.add_argument("--k2", type=int, default=6)
    ap.add_argument("--lam", type=float, default=0.3)
    ap.add_argument("--wide", type=int, default=64, help="neighbours kept per row by the N x N sweep")
    ap.add_argument("--ids", type=int, default=50_000, help="identities (gallery draws from 1.2x as many), 32 cameras")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    engine = importlib.import_module("multimodal-reid_b200.engine")
    eng = engine.Engine(dev)
    cm = eng.comm
    q, g, d, K = args.q, args.g, args.d, args.topk
    gen = torch.Generator(device=dev).manual_seed(0)
    # SURVEY 8(d) generator: identity centroids ~ N(0, I), sample = centroid + 2.5 * N(0, I); the gallery draws from
    # 1.2x as many identities (distractors), cameras uniform over 32
    n_gid = int(args.ids * 1.2)
    qp = torch.randint(0, args.ids, (q,), generator=gen, device=dev)
    gp = torch.randint(0, n_gid, (g,), generator=gen, device=dev)
    qc = torch.randint(0, 32, (q,), generator=gen, device=dev)
    gc = torch.randint(0, 32, (g,), generator=gen, device=dev)
    cent = torch.randn(n_gid, d, generator=gen, device=dev)
    pid_all = torch.cat([qp, gp])
    feats = torch.empty((q + g, d), dtype=torch.float32, device=dev)
    for a in range(0, q + g, 1 << 18):               # bounded temporaries
        b = min(q + g, a + (1 << 18))
        feats[a:b] = cent[pid_all[a:b]] + 2.5 * torch.randn(b - a, d, generator=gen, device=dev)
    del cent, pid_all
    allf = eng.prepare(feats, True)
    del feats
    Q, G = allf.rows(slice(cm.rank, q, cm.world)), allf.rows(slice(q, q + g))

    mine = slice(cm.rank, q, cm.world)
    qp_l, qc_l = qp[mine].contiguous(), qc[mine].contiguous()
    score_out = {}

    pids_all, cams_all = torch.cat([qp, gp]), torch.cat([qc, gc])
    diag_out = {}

    def step():
        if args.rerank:
            (cmc, mAP, nv, _, _), (idx, val), diag = eng.evaluate_streamed_reranked(
                allf, q, pids_all, cams_all, args.k1, args.k2, args.lam, topk=K, Kw=args.wide,
                rows_per_pass=args.rows_per_pass, cols_per_pass=args.cols_per_pass)
            score_out.update(cmc=cmc, mAP=mAP, nv=nv)
            diag_out.update(diag)
        elif args.score:
            (cmc, mAP, nv, _, _), (idx, val) = eng.evaluate_streamed(Q, G, qp_l, gp, qc_l, gc, 50, K,
                                                                     rows_per_pass=args.rows_per_pass,
                                                                     cols_per_pass=args.cols_per_pass, q_total=q)
            score_out.update(cmc=cmc, mAP=mAP, nv=nv)
        else:
            idx, val = eng.retrieve_topk(Q, G, K, rows_per_pass=args.rows_per_pass, cols_per_pass=args.cols_per_pass)
        if cm.world > 1:
            idx, val = cm.gather_rows(idx, q), cm.gather_rows(val, q)
        return idx, val

    for _ in range(args.warmup):
        step()
    eng.enable_timers(True)
    n0 = eng.lib.reid_launch_count()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        idx, val = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item()) / args.steps
    launches = int(eng.lib.reid_launch_count() - n0)
    stages = {k: v / args.steps for k, v in eng.stage_ms().items()}
    if rank == 0:
        peaks = {}
        pth = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pth):
            with open(pth) as f:
                peaks = json.load(f)
        # the sustained figure for contraction stages that run for seconds at the power cap; short per-rank stages (8 ranks)
        # run at burst clocks, where the burst figure is the honest denominator
        bf16_sus = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
        bf16_burst = float(peaks.get("bf16_tflops", bf16_sus))
        hbm = float(peaks.get("hbm_gbs", 6450.0))
        rows_local = cm.local_count(q)
        flops = 2.0 * rows_local * g * d
        dist_ms = stages.get("dist", float("nan"))
        # one q x g sweep in every mode (the match distances of the scoring modes are contracted as gathered pairs, stage
        # "pairs", not counted here)
        if args.rerank:                          # + this rank's share of the N x N sweep (half of it from the upper triangle)
            half = diag_out.get("nn_sweep") == "upper half"
            flops += 2.0 * (q + g) * (q + g) * d / world * (0.5 if half else 1.0)
        issued = 3.0 * flops / (dist_ms * 1e-3) / 1e12
        long_run = dist_ms > 1000.0
        bf16 = bf16_sus if long_run else bf16_burst
        line = {"metric": ("k-reciprocal re-ranked top-k + CMC/mAP" if args.rerank else
                           "top-k retrieval + CMC/mAP" if args.score else "top-k retrieval")
                          + " q x g pairs/sec (matrix never materialised)",
                "value": q * g / (ms * 1e-3), "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
                "dtype": "f32 (fp16-split tensor-core contraction)", "data": "synthetic",
                "config": {"workload": "scaled retrieval", "q": q, "g": g, "d": d, "topk": K,
                           "rows_per_pass": args.rows_per_pass, "cols_per_pass": args.cols_per_pass,
                           "rerank": ({"k1": args.k1, "k2": args.k2, "lambda": args.lam, "wide": args.wide, **diag_out}
                                      if args.rerank else None),
                           "inputs": "larger than L2 (no flush needed)"},
                "gpu_launches": launches, "stages_ms": stages,
                # rectangular panels: every tile is computed (no symmetry to exploit), three fp16 MMAs per K step; tiles are
                # padded to 256 x 256, which the panel shapes used here divide evenly
                "roofline": {"kernel": "dist_tc2_kernel (rectangular panels)", "bound": "tensor",
                             "achieved": issued, "peak": bf16, "unit": "TFLOP/s", "frac": issued / bf16, "ms": dist_ms,
                             "achieved_is": "issued fp16 MMA flops (3 x algorithmic 2*rows*cols*d) / CUDA-event time of the "
                                            "contraction launches",
                             "algorithmic_tflops": flops / (dist_ms * 1e-3) / 1e12,
                             "peak_source": "MEASURED_PEAKS.json bf16 " + ("sustained (contraction stages > 1 s)" if long_run
                                                                            else "burst (contraction stages < 1 s)")},
                "hbm_panel": {"bytes_written_per_step": 4.0 * rows_local * g * (1 + 1 / 32),
                              "GB/s_over_dist": 4.0 * rows_local * g * (1 + 1 / 32) / (dist_ms * 1e-3) / 1e9,
                              "peak": hbm},
                "score": ({"mAP": float(score_out["mAP"].item()), "rank1": float(score_out["cmc"][0].item()),
                           "n_valid": int(score_out["nv"].item()), "sweeps": 2 if args.rerank else 1}
                          if (args.score or args.rerank) else None),
                "first_row": {"idx": idx[0, :5].tolist(), "dist": [round(float(x), 6) for x in val[0, :5]]}}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
