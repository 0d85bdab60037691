"""Deterministic synthetic ReID features.

Identities get Gaussian centroids, samples are centroid + sigma * noise, gallery ids are drawn
from 1.2x the query identity pool (20 % distractors), cameras are uniform.  Everything comes from
a seeded CPU ``torch.Generator`` so the same tensors appear in this container and on the GPU box.

The named shapes follow the dataset docstrings the reference cites
(datasets/dataset_market.py:15, dataset_dukemtmc.py:24, dataset_veri.py:19, dataset_msmt17.py:20).
"""
from __future__ import annotations

import dataclasses

import torch


@dataclasses.dataclass(frozen=True)
class Shape:
    name: str
    q: int
    g: int
    d: int
    ids: int
    cams: int
    k1: int = 20
    k2: int = 6
    lam: float = 0.3


SHAPES = {
    "market": Shape("market", 3368, 15913, 512, 751, 6),
    "duke": Shape("duke", 2228, 17661, 512, 702, 8),
    "veri": Shape("veri", 1678, 11579, 768, 200, 20),
    "msmt17": Shape("msmt17", 11659, 82161, 512, 3060, 15),
    # down-scaled instances for CPU-side parity (same generator, same density of positives)
    "tiny": Shape("tiny", 60, 300, 64, 20, 4, k1=20, k2=6),
    "small": Shape("small", 200, 1200, 128, 60, 6, k1=20, k2=6),
    "mid": Shape("mid", 700, 4300, 512, 160, 6, k1=20, k2=6),
}


def scaled(shape: Shape, q: int, g: int) -> Shape:
    """Same generator parameters at a reduced (q, g); identities shrink with the gallery."""
    ids = max(4, int(round(shape.ids * g / shape.g)))
    return dataclasses.replace(shape, name=f"{shape.name}@{q}x{g}", q=q, g=g, ids=ids)


def make(shape: Shape, seed: int = 0, sigma: float | None = None):
    """Returns (qf, gf, q_pids, g_pids, q_cams, g_cams); features fp32 un-normalised, ids int64."""
    gen = torch.Generator().manual_seed(seed)
    d = shape.d
    if sigma is None:
        sigma = 2.5  # non-saturated task at d=512 (survey probe: mAP 0.45 -> 0.92 re-ranked)
    n_ids_all = int(round(shape.ids * 1.2))
    cent = torch.randn(n_ids_all, d, generator=gen)
    q_pids = torch.randint(0, shape.ids, (shape.q,), generator=gen, dtype=torch.int64)
    g_pids = torch.randint(0, n_ids_all, (shape.g,), generator=gen, dtype=torch.int64)
    q_cams = torch.randint(0, shape.cams, (shape.q,), generator=gen, dtype=torch.int64)
    g_cams = torch.randint(0, shape.cams, (shape.g,), generator=gen, dtype=torch.int64)
    qf = cent[q_pids] + sigma * torch.randn(shape.q, d, generator=gen)
    gf = cent[g_pids] + sigma * torch.randn(shape.g, d, generator=gen)
    return qf, gf, q_pids, g_pids, q_cams, g_cams
