"""Drop-in for the distance / hard-mining part of the reference's ``losses.py``.

``euclidean_dist`` (losses.py:14-28), ``hard_example_mining`` (:31-83) and the two triplet losses built on
them (``WeightedRegularizedTriplet`` :86-108, ``WeightedRegularizedTripletXBM`` :111-152) keep their names,
arguments and return values; forward and backward run as sm_100a kernels of libreid_b200.so (train.cu).
torch is used for device buffers, autograd bookkeeping and the final N-element ranking loss, as in the
reference.  No CPU path: CPU tensors raise."""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.nn import functional as F

from . import _lib
from .engine import _ptr, _stream


def _dev_f32(t: torch.Tensor, what: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.ReidError(f"{what} must be a CUDA tensor: reid_b200 has no CPU path")
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.stride(-1) == 1 else t.contiguous()


def _labels(t: torch.Tensor, device) -> torch.Tensor:
    return t.to(device=device, dtype=torch.int64).contiguous()


TC_MIN_ROWS = 1024          # from this batch size on the forward is a real GEMM: tensor cores (reid_pair_dist_tc)


def _pair_dist_forward(xs: torch.Tensor, ys: torch.Tensor, same: bool, s) -> torch.Tensor:
    """losses.py:14-28 into a fresh [m, n] tensor (a view of a pitched buffer only when n % 4 != 0 on the tensor path).  Batches below TC_MIN_ROWS run the
    float32 CUDA-core kernel (latency-bound, one launch); larger ones the tcgen05 contraction of the evaluation path.
    `same`: x and y are the same tensor object (upper triangle + mirrored stores)."""
    m, d = xs.shape
    n = ys.shape[0]
    tc = min(m, n) >= TC_MIN_ROWS
    ld = (n + 3) // 4 * 4 if tc else n          # the tensor-core store needs 16-byte aligned rows; contiguous otherwise
    out = torch.empty((m, ld), dtype=torch.float32, device=xs.device)
    if tc:
        y_arg = xs if same else ys
        nbytes = int(_lib.load().reid_pair_dist_tc_workspace_bytes(m, n, d, 1 if same else 0))
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=xs.device)
        _lib.call("reid_pair_dist_tc", _ptr(xs), xs.stride(0), m, _ptr(y_arg), y_arg.stride(0), n, d, _ptr(ws), nbytes,
                  _ptr(out), ld, s)
    else:
        _lib.call("reid_pair_dist", _ptr(xs), xs.stride(0), m, _ptr(ys), ys.stride(0), n, d, _ptr(out), ld, s)
    return out[:, :n]


class _PairDist(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y):
        xs, ys = _dev_f32(x, "x"), _dev_f32(y, "y")
        m, d = xs.shape
        n = ys.shape[0]
        if ys.shape[1] != d:
            raise ValueError(f"feature widths differ: {d} vs {ys.shape[1]}")
        with torch.cuda.device(xs.device):
            out = _pair_dist_forward(xs, ys, x is y, _stream())
        ctx.save_for_backward(xs, ys, out)
        ctx.dtypes = (x.dtype, y.dtype)
        return out

    @staticmethod
    def backward(ctx, G):
        xs, ys, dist = ctx.saved_tensors
        G = _dev_f32(G, "grad").contiguous()
        m, d = xs.shape
        n = ys.shape[0]
        need_x, need_y = ctx.needs_input_grad
        gx = torch.empty_like(xs, memory_format=torch.contiguous_format) if need_x else None
        gy = torch.empty_like(ys, memory_format=torch.contiguous_format) if need_y else None
        with torch.cuda.device(xs.device):
            _lib.call("reid_pair_dist_grad", _ptr(xs), xs.stride(0), m, _ptr(ys), ys.stride(0), n, d, _ptr(dist),
                      dist.stride(0), _ptr(G), n, _ptr(gx), d, _ptr(gy), d, 0, _stream())
        gx = gx.to(ctx.dtypes[0]) if gx is not None else None
        gy = gy.to(ctx.dtypes[1]) if gy is not None else None
        return gx, gy


def euclidean_dist(x, y):
    """losses.py:14-28 — [m, d], [n, d] -> [m, n] Euclidean distances (sqrt of the clamped squared form)."""
    return _PairDist.apply(x, y)


class _HardMine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dist_mat, lab_r, lab_c, skip_diag):
        D = _dev_f32(dist_mat, "dist_mat")
        m, n = D.shape
        dev = D.device
        ap = torch.empty(m, dtype=torch.float32, device=dev)
        an = torch.empty(m, dtype=torch.float32, device=dev)
        pi = torch.empty(m, dtype=torch.int64, device=dev)
        ni = torch.empty(m, dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            _lib.call("reid_hard_mine", _ptr(D), D.stride(0), m, n, _ptr(lab_r), _ptr(lab_c), int(skip_diag), _ptr(ap),
                      _ptr(an), _ptr(pi), _ptr(ni), _stream())
        ctx.save_for_backward(pi, ni)
        ctx.shape = (m, n)
        ctx.mark_non_differentiable(pi, ni)
        return ap, an, pi, ni

    @staticmethod
    def backward(ctx, g_ap, g_an, _gp, _gn):
        pi, ni = ctx.saved_tensors
        m, n = ctx.shape
        g_ap = g_ap.contiguous().float()
        g_an = g_an.contiguous().float()
        G = torch.empty((m, n), dtype=torch.float32, device=pi.device)
        with torch.cuda.device(pi.device):
            _lib.call("reid_hard_mine_grad", _ptr(g_ap), _ptr(g_an), _ptr(pi), _ptr(ni), m, n, _ptr(G), n, _stream())
        return G, None, None, None


def hard_example_mining(dist_mat, labels, return_inds=False):
    """losses.py:31-83 — hardest positive (farthest same-label sample, the anchor itself included) and hardest
    negative (nearest other-label sample) of every anchor.  Unlike the reference, identities need not have
    the same number of samples (its ``view(N, -1)`` requires that)."""
    assert len(dist_mat.size()) == 2
    assert dist_mat.size(0) == dist_mat.size(1)
    lab = _labels(labels, dist_mat.device)
    dist_ap, dist_an, p_inds, n_inds = _HardMine.apply(dist_mat, lab, lab, False)
    if return_inds:
        return dist_ap, dist_an, p_inds, n_inds
    return dist_ap, dist_an


class _MinedPairs(torch.autograd.Function):
    """euclidean_dist + hard mining as ONE autograd node for the losses that only use the mined pairs
    (losses.py:96-98, :121-139): forward = reid_pair_dist + reid_hard_mine, backward = reid_triplet_grad
    (O((m + n) d), no m x n gradient matrix)."""

    @staticmethod
    def forward(ctx, x, y, lab_r, lab_c, skip_diag):
        xs, ys = _dev_f32(x, "x"), _dev_f32(y, "y")
        m, d = xs.shape
        n = ys.shape[0]
        dev = xs.device
        ap = torch.empty(m, dtype=torch.float32, device=dev)
        an = torch.empty(m, dtype=torch.float32, device=dev)
        pi = torch.empty(m, dtype=torch.int64, device=dev)
        ni = torch.empty(m, dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            s = _stream()
            dist = _pair_dist_forward(xs, ys, x is y, s)
            _lib.call("reid_hard_mine", _ptr(dist), dist.stride(0), m, n, _ptr(lab_r), _ptr(lab_c), int(skip_diag), _ptr(ap),
                      _ptr(an), _ptr(pi), _ptr(ni), s)
        ctx.save_for_backward(xs, ys, ap, an, pi, ni)
        ctx.dtypes = (x.dtype, y.dtype)
        return ap, an

    @staticmethod
    def backward(ctx, g_ap, g_an):
        xs, ys, ap, an, pi, ni = ctx.saved_tensors
        m, d = xs.shape
        n = ys.shape[0]
        need_x, need_y = ctx.needs_input_grad[:2]
        g_ap = g_ap.contiguous().float()
        g_an = g_an.contiguous().float()
        gx = torch.empty((m, d), dtype=torch.float32, device=xs.device) if need_x else None
        gy = torch.empty((n, d), dtype=torch.float32, device=xs.device) if need_y else None
        with torch.cuda.device(xs.device):
            _lib.call("reid_triplet_grad", _ptr(xs), xs.stride(0), m, _ptr(ys), ys.stride(0), n, d, _ptr(ap), _ptr(an),
                      _ptr(pi), _ptr(ni), _ptr(g_ap), _ptr(g_an), _ptr(gx), d, _ptr(gy), d, 0, _stream())
        gx = gx.to(ctx.dtypes[0]) if gx is not None else None
        gy = gy.to(ctx.dtypes[1]) if gy is not None else None
        return gx, gy, None, None, None


class WeightedRegularizedTriplet(object):
    """losses.py:86-108 (margin ranking loss, or soft-margin when ``margin`` is None)."""

    def __init__(self, margin=None):
        if margin is not None:
            self.ranking_loss = nn.MarginRankingLoss(margin=margin)
        else:
            self.ranking_loss = nn.SoftMarginLoss()
        self.margin = margin

    def _rank(self, dist_ap, dist_an):
        y = torch.ones_like(dist_an)
        if self.margin is not None:
            return self.ranking_loss(dist_an, dist_ap, y)
        return self.ranking_loss(dist_an - dist_ap, y)

    def __call__(self, global_feat, labels, normalize_feature=False):
        if normalize_feature:
            global_feat = F.normalize(global_feat, dim=-1)
        lab = _labels(labels, global_feat.device)
        dist_ap, dist_an = _MinedPairs.apply(global_feat, global_feat, lab, lab, False)
        return self._rank(dist_ap, dist_an)


class WeightedRegularizedTripletXBM(WeightedRegularizedTriplet):
    """losses.py:111-152 with the masks the code sets up (:126-128): positives = same label in the memory
    bank minus the anchor's own slot (the first N bank entries are the batch), negatives = different label.

    The reference then indexes ``dist_mat[is_pos]`` with a *long* 0/1 tensor (:132, :137).  That is integer (advanced)
    indexing, not masking: it gathers rows 0 and 1 of the matrix for every element of the mask, so after ``view(N, -1)``
    every anchor gets dist_ap = max(dist_mat[0:2]) and dist_an = min(dist_mat[0:2]) (each mask row holds both values).
    Default: the intended masks.  ``reference_compat=True`` reproduces what the reference returns, value and gradient
    (tests/test_gpu_losses.py compares it with the unmodified reference on the GPU box), so that a run can be
    continued bit-compatibly; INTEGRATION.md 2b states the difference."""

    def __init__(self, margin=None, reference_compat: bool = False):
        super().__init__(margin)
        self.reference_compat = reference_compat

    def __call__(self, global_feat, labels, global_feat_xbm, labels_xbm, normalize_feature=False):
        if normalize_feature:
            global_feat = F.normalize(global_feat, dim=-1)
            global_feat_xbm = F.normalize(global_feat_xbm, dim=-1)
        dev = global_feat.device
        if self.reference_compat:
            n = global_feat.shape[0]
            lab, lab_x = _labels(labels, dev), _labels(labels_xbm, dev)
            is_pos = lab[:, None].eq(lab_x[None, :]).long()
            is_pos[:, :n] -= torch.eye(n, dtype=torch.long, device=dev)
            rows = euclidean_dist(global_feat[:2], global_feat_xbm)            # the only rows the indexing ever reads
            # a mask row that is all 0 (all 1) gathers row 0 (row 1) only; with both values present it gathers both
            def pick(mask):
                has1, has0 = mask.bool().any(1), (~mask.bool()).any(1)
                r0 = rows[0].max() if mask is is_pos else rows[0].min()
                r1 = rows[1].max() if mask is is_pos else rows[1].min()
                both = torch.maximum(r0, r1) if mask is is_pos else torch.minimum(r0, r1)
                return torch.where(has1 & has0, both, torch.where(has1, r1, r0))
            return self._rank(pick(is_pos), pick(1 - is_pos))
        dist_ap, dist_an = _MinedPairs.apply(global_feat, global_feat_xbm, _labels(labels, dev),
                                             _labels(labels_xbm, dev), True)
        return self._rank(dist_ap, dist_an)
