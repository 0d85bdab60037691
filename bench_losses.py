"""Timing of the training-side distance ops: triplet loss forward + backward on one B200.

    python bench_losses.py [--steps 200]

Prints one JSON line per batch shape: ms per forward+backward through the drop-in (losses.WeightedRegularizedTriplet ->
reid_pair_dist / reid_hard_mine / reid_hard_mine_grad / reid_pair_dist_grad), next to a torch-eager restatement of the
same math on the same GPU (what the reference's losses.py:14-108 executes when its tensors live on the device: ~25 small
library kernels).  These ops are latency-bound (B x B matrices of KBs), so the figure of merit is time per step, and the
roofline is the launch floor, not bandwidth: stated in DESIGN.md §4.5."""
import argparse
import json

import torch


def torch_eager_triplet(x, labels, margin=0.3):
    m = x.size(0)
    xx = torch.pow(x, 2).sum(1, keepdim=True).expand(m, m)
    dist = (xx + xx.t() - 2 * torch.matmul(x, x.t())).clamp(min=1e-12).sqrt()
    pos = labels.expand(m, m).eq(labels.expand(m, m).t())
    ap = torch.max(dist[pos].contiguous().view(m, -1), 1)[0]
    an = torch.min(dist[~pos].contiguous().view(m, -1), 1)[0]
    return torch.nn.functional.margin_ranking_loss(an, ap, torch.ones_like(an), margin=margin)


def timed(fn, steps):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    args = ap.parse_args()
    import reid_losses as losses
    import importlib
    lib = importlib.import_module("multimodal-reid_b200._lib").load()
    torch.backends.cuda.matmul.allow_tf32 = False
    for B, K, d in ((64, 4, 512), (64, 4, 1280), (256, 4, 1280), (512, 8, 1280), (1024, 8, 3072), (2048, 8, 3072),
                    (4096, 8, 2048)):
        g = torch.Generator().manual_seed(0)
        labels = torch.arange(B // K).repeat_interleave(K).cuda()
        x0 = torch.randn(B, d, generator=g).cuda()
        crit = losses.WeightedRegularizedTriplet(0.3)

        def ours():
            x = x0.detach().requires_grad_(True)
            crit(x, labels).backward()
            return x.grad

        def eager():
            x = x0.detach().requires_grad_(True)
            torch_eager_triplet(x, labels).backward()
            return x.grad

        ga, gb = ours(), eager()
        n0 = lib.reid_launch_count()
        ours()
        launches = lib.reid_launch_count() - n0
        err = (ga - gb).abs().max().item() / max(gb.abs().max().item(), 1e-30)
        t_ours, t_eager = timed(ours, args.steps), timed(eager, args.steps)
        print(json.dumps({"metric": "triplet loss fwd+bwd", "B": B, "d": d, "ms_per_step": t_ours,
                          "torch_eager_ms_per_step": t_eager, "gpu_launches": int(launches),
                          "max_rel_grad_diff_vs_eager": err,
                          "forward": "tcgen05 fp16 hi/lo" if B >= 1024 else "f32 CUDA cores", "dtype": "f32"}))


if __name__ == "__main__":
    main()
