"""Freezes outputs and autograd gradients of the UNMODIFIED reference losses.py (imported from /root/reference,
never copied) into tests/golden/losses/*.npz.  TEST INFRASTRUCTURE ONLY.  Run in the build container:
    python oracle/make_golden_losses.py
Per case: features x [B, d] (identities with equal sample counts, as losses.py:44-45 requires), labels;
  euclidean_dist(x, x), euclidean_dist(x, y)                                  losses.py:14-28
  hard_example_mining(dist, labels, return_inds=True)                         losses.py:31-83
  WeightedRegularizedTriplet(margin)(x, labels[, normalize_feature]) and its gradient w.r.t. x   losses.py:86-108
  d sum(dist * R) / dx, dy for a fixed random R (dense autograd of euclidean_dist)
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("REID_REFERENCE_ROOT", "/root/reference")
CASES = [  # name, identities, samples per identity, d, sigma, seed
    ("b32_d64", 8, 4, 64, 1.0, 0),
    ("b64_d512", 16, 4, 512, 2.5, 1),
    ("b96_d1280", 24, 4, 1280, 2.5, 2),
    ("b60_d77", 12, 5, 77, 0.5, 3),
]


def make(P, K, d, sigma, seed):
    g = torch.Generator().manual_seed(seed)
    cent = torch.randn(P, d, generator=g)
    labels = torch.arange(P).repeat_interleave(K)
    perm = torch.randperm(P * K, generator=g)
    labels = labels[perm] * 7 + 3                      # arbitrary int64 ids
    x = cent[(labels - 3) // 7] + sigma * torch.randn(P * K, d, generator=g)
    y = torch.randn(P * K + 9, d, generator=g)
    R = torch.rand(P * K, P * K + 9, generator=g) - 0.3
    return x, labels, y, R


def main():
    spec = importlib.util.spec_from_file_location("_ref_losses", os.path.join(REF, "losses.py"))
    L = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(L)
    out_dir = os.path.join(ROOT, "tests", "golden", "losses")
    os.makedirs(out_dir, exist_ok=True)
    for name, P, K, d, sigma, seed in CASES:
        x, labels, y, R = make(P, K, d, sigma, seed)
        rec = dict(x=x.numpy(), labels=labels.numpy(), y=y.numpy(), R=R.numpy())
        with torch.no_grad():
            dxx = L.euclidean_dist(x, x)
            rec["dist_xx"] = dxx.numpy()
            rec["dist_xy"] = L.euclidean_dist(x, y).numpy()
            ap, an, pi, ni = L.hard_example_mining(dxx, labels, return_inds=True)
            rec.update(dist_ap=ap.numpy(), dist_an=an.numpy(), p_inds=pi.numpy(), n_inds=ni.numpy())
        for tag, margin, norm in (("m03", 0.3, False), ("soft", None, False), ("m03n", 0.3, True)):
            xv = x.clone().requires_grad_(True)
            loss = L.WeightedRegularizedTriplet(margin)(xv, labels, normalize_feature=norm)
            loss.backward()
            rec[f"loss_{tag}"] = loss.detach().numpy()
            rec[f"grad_{tag}"] = xv.grad.numpy()
        xv = x.clone().requires_grad_(True)
        yv = y.clone().requires_grad_(True)
        (L.euclidean_dist(xv, yv) * R).sum().backward()
        rec["gx_dense"] = xv.grad.numpy()
        rec["gy_dense"] = yv.grad.numpy()
        path = os.path.join(out_dir, f"{name}.npz")
        np.savez_compressed(path, **rec)
        print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
