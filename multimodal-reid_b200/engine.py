"""Device-side orchestration of the retrieval-evaluation hot path.

torch supplies device memory, the current stream and (for N > 1 ranks) the NCCL collectives; every
arithmetic step is a kernel in libreid_b200.so reached through the C-ABI (include/reid_b200.h).

Multi-GPU layout: rank r of P owns the probe rows {i : i mod P == r}.  Queries are
the first rows, so both the N probes (stages K3–K7) and the q queries (K9–K12) are spread evenly.
Three all-gathers move the neighbour lists, the V rows and the V_qe rows (MB-sized); the CMC
histogram is an all-reduce and the per-query APs an all-gather.  Every row is produced by exactly
one rank with the same kernel, so results do not depend on P.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os

import numpy as np
import torch

from . import _lib
from . import fabric as _fabric

METRIC_SQEUCLIDEAN = 0
METRIC_ARCCOS = 1
ENGINES = {"auto": 0, "simt": 1, "tcgen05": 2}
GROUP_MAX = 8192             # same-identity gallery samples per query the streamed scorers hold (csrc/stream_score.cu)


def _ptr(t):
    return C.c_void_p(0) if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def half_k(k1: int) -> int:
    return _lib.load().reid_half_k(int(k1))


def v_row_capacity(k1: int) -> int:
    return _lib.load().reid_v_row_capacity(int(k1))


_ONE_BITS = 0x3F800000      # float32 1.0


@dataclasses.dataclass
class Features:
    """Prepared feature rows (engine.prepare): float32 rows padded to dp, their float16 hi/lo split
    for the tensor cores, squared norms, norms, and the matrix-wide max|x| that fixes the split's
    power-of-two scale.  Row subsets are strided views of the same buffers (no copies)."""
    x32: torch.Tensor      # [n, dp] float32 (view; row pitch = x32.stride(0))
    hi: torch.Tensor       # [n, dp] float16
    lo: torch.Tensor       # [n, dp] float16
    sq: torch.Tensor       # [n] float32
    nrm: torch.Tensor      # [n] float32
    amax: torch.Tensor     # [1] int32 (float bits)
    d: int

    @property
    def n(self) -> int:
        return self.hi.shape[0]

    @property
    def dp(self) -> int:
        return self.hi.shape[1]

    def rows(self, sl: slice) -> "Features":
        """Row subset; sq/nrm are compacted because the kernels index them by local row."""
        return Features(self.x32[sl], self.hi[sl], self.lo[sl], self.sq[sl].contiguous(), self.nrm[sl].contiguous(),
                        self.amax, self.d)


class Comm:
    """Cyclic row ownership + the collectives the path needs.  world == 1 is the identity."""

    def __init__(self, group=None, single: bool = False):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        if single:
            self.rank, self.world = 0, 1
        elif dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1

    def local_count(self, n: int) -> int:
        """#{i < n : i mod world == rank}"""
        return (n - self.rank + self.world - 1) // self.world if n > self.rank else 0

    def _all_gather(self, local: torch.Tensor) -> torch.Tensor:
        """One all-gather of equally shaped contiguous tensors -> [world, *shape].  The output is allocated in the
        concatenated form (world * shape[0], ...), which every backend accepts (gloo rejects the stacked form)."""
        flat = local.reshape(1) if local.dim() == 0 else local
        out = flat.new_empty((self.world * flat.shape[0],) + tuple(flat.shape[1:]))
        self.dist.all_gather_into_tensor(out, flat, group=self.group)
        return out.view((self.world,) + tuple(local.shape))

    def gather_rows(self, local: torch.Tensor, n: int) -> torch.Tensor:
        """local[l] is global row rank + l*world; returns all n rows in global order."""
        if self.world == 1:
            return local
        nmax = (n + self.world - 1) // self.world
        pad = local
        if local.shape[0] != nmax:
            pad = local.new_zeros((nmax,) + tuple(local.shape[1:]))
            pad[: local.shape[0]] = local
        out = self._all_gather(pad.contiguous())
        # [P, nmax, ...] -> interleave -> [nmax*P, ...]
        out = out.transpose(0, 1).reshape((nmax * self.world,) + tuple(local.shape[1:]))
        return out[:n].contiguous()

    def gather_flat(self, local: torch.Tensor, size_max: int) -> torch.Tensor:
        """Variable-length 1-D pools padded to size_max; returns [world, size_max]."""
        if self.world == 1:
            return local.reshape(1, -1)
        pad = local.new_zeros((size_max,))
        pad[: local.shape[0]] = local
        return self._all_gather(pad)

    def gather_stack(self, local: torch.Tensor) -> torch.Tensor:
        """All ranks pass the same shape; returns [world, *shape] (rank-major, one NCCL call)."""
        local = local.contiguous()
        if self.world == 1:
            return local.unsqueeze(0)
        return self._all_gather(local)

    def nmax(self, n: int) -> int:
        return (n + self.world - 1) // self.world

    def block_bounds(self, n: int, rank: int | None = None) -> tuple[int, int]:
        """Contiguous block [lo, hi) of n rows that rank uploads in upload_sharded (ceil(n / world) rows each)."""
        r = self.rank if rank is None else rank
        nm = self.nmax(n)
        return min(n, r * nm), min(n, (r + 1) * nm)

    def upload_sharded(self, parts, device) -> torch.Tensor:
        """Host feature batches (the list R1_mAP_eval.update() collected, evaluate.py:104-106) -> the full [N, d]
        matrix on every rank, with each rank copying only ITS contiguous 1/world block over PCIe and the blocks
        all-gathered over NVLink: host->device bytes per rank fall by the factor world (at 8 ranks the replicated
        upload of MSMT17's 192 MB shared one host's PCIe/DRAM bandwidth eight ways)."""
        n = sum(int(f.shape[0]) for f in parts)
        d = int(parts[0].shape[1])
        nm = self.nmax(n)
        lo, hi = self.block_bounds(n)
        mine = torch.empty((nm, d), dtype=torch.float32, device=device)
        off = 0
        for f in parts:
            rows = int(f.shape[0])
            a, b = max(lo, off), min(hi, off + rows)
            if a < b:
                mine[a - lo: b - lo].copy_(f[a - off: b - off], non_blocking=True)
            off += rows
        if hi - lo < nm:
            mine[hi - lo:].zero_()
        if self.world == 1:
            return mine[:n]
        return self._all_gather(mine).view(self.world * nm, d)[:n]

    def pad_rows(self, local: torch.Tensor, n: int) -> torch.Tensor:
        """Zero-pads this rank's rows to ceil(n / world) so that every rank contributes the same shape."""
        nm = self.nmax(n)
        if local.shape[0] == nm:
            return local
        pad = local.new_zeros((nm,) + tuple(local.shape[1:]))
        pad[: local.shape[0]] = local
        return pad

    @staticmethod
    def interleave(stacked: torch.Tensor, n: int) -> torch.Tensor:
        """stacked[r, l] is global row r + l*world -> [n, ...] in global order (a view; copy with .contiguous())."""
        P, nm = stacked.shape[0], stacked.shape[1]
        return stacked.transpose(0, 1).reshape((nm * P,) + tuple(stacked.shape[2:]))[:n]

    @staticmethod
    def pool_layout(ub: torch.Tensor, P: int):
        """ub: [nm * P] upper bounds of the rows' union sizes in GLOBAL row order (zero for padding rows).
        Rank p stores its rows' unions back to back in its own pool, in row order.  Returns (totals[P], excl[P, nm]):
        pool sizes and the offset of row p + l*P inside rank p's pool -- what every rank can compute without
        talking to the others."""
        nm = ub.shape[0] // P
        ub_rm = ub.view(nm, P).t()                               # [P, nm]: row l of rank p
        incl = torch.cumsum(ub_rm, dim=1, dtype=torch.int64)
        return incl[:, -1], incl - ub_rm

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            self.dist.all_reduce(t, group=self.group)
        return t

    def max_int(self, v: int, device) -> int:
        if self.world == 1:
            return v
        t = torch.tensor([v], dtype=torch.int64, device=device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return int(t.item())


class Engine:
    def __init__(self, device=None, engine: str = "auto", comm: Comm | None = None):
        if not torch.cuda.is_available():
            raise _lib.ReidError("reid_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.lib = _lib.load()
        rc = self.lib.reid_device_check(self.device.index or 0)
        if rc != 0:
            raise _lib.ReidError(self.lib.reid_last_error().decode())
        self.engine = ENGINES[engine]
        self.comm = comm or Comm()
        self.timers = None          # dict stage -> list[(start_event, end_event)] when profiling
        self.nvtx = bool(int(__import__("os").environ.get("REID_NVTX", "0")))   # stage spans as NVTX ranges (nsys / ncu --nvtx)
        # Jaccard stage: packed per-query overlap records (jaccard_sparse.cu) with up to `sparse_cap` touched probes;
        # denser queries fall back to the dense shared-memory row (jaccard_score.cu).  sparse_cap = 0: dense only.
        self.sparse_cap = 8192
        # N x N contraction of one rank owning every row: upper-triangle tiles only, mirrored (dist_tc2.cu).
        # Ranks of a multi-GPU job own strided row subsets and always run the full schedule.
        self.symmetric = True
        self.upper_only = True      # ... and, when nobody needs the full matrix, without storing the mirrored half
        self.upper_only_min_rows = 0
        self.stream_h2d = True              # evaluator: copy host features chunk by chunk under the contraction
        self.shard_h2d = True               # evaluator on several ranks: each uploads 1/world of the host rows, all-gather over NVLink
        self._copy_stream = None
        self.scatter_select = True          # upper-only selection in scatter form (reid_topk_rows_scatter)
        self.c_entry = True                 # one rank: evaluate_reranked goes through the single C entry (reid_eval_reranked)
        self._c_ws = None
        # ... replayed as ONE CUDA-graph launch for problems of fewer rows (0: never; REID_GRAPH_BELOW overrides)
        self.graph_below = int(__import__("os").environ.get("REID_GRAPH_BELOW", "50000"))
        self._graphs = {}
        self._graph_last_key, self._graph_run = None, 0
        self._overlap_hint = {}             # (N, q, k1, k2) -> largest overlap list of the previous streamed re-ranking
        self._capturing = False
        self.qe_nosync_bytes = 4 << 30       # budget for sizing the query-expansion pools without a host read
        self._ubmax_pinned = torch.zeros(1, dtype=torch.int64).pin_memory()
        self._ubmax_event = torch.cuda.Event()
        self._ubmax_pending = False
        self._ubmax_seen = 256
        self.qe_replicate_below = 300_000   # N * k2 up to which every rank runs the query expansion for all rows
        self._cw_hint = {}          # k1 -> width of the V rows exchanged between ranks (see rerank_from_panel)
        # world > 1: symmetric contraction sharded by tile rows + exchanges through peer memory (fabric.py); the NCCL
        # all-gather path of round 1 remains as the fallback when CUDA IPC / peer access is unavailable
        self.use_fabric = True
        self.sym_sweep = os.environ.get("REID_SYM_SWEEP", "1") != "0"    # streamed re-ranking: N x N lists from half the matrix
        self.sym_cap = 512          # candidate records per column and push of that sweep (Kw + sym_cap <= 1024)
        self.fabric = None
        self._fabric_failed = False
        self.fabric_pool_bytes = 24 << 30   # largest query-expansion pool the arena may hold (structural bound)
        self._layouts, self._tile_cache, self._own_cache = {}, {}, {}
        self.last_tiles = 0
        self.last_path = "single"           # which multi-rank path the last rerank_features call took: fabric | nccl | single

    # ------------------------------------------------------------------ timing helpers
    def enable_timers(self, on=True):
        self.timers = {} if on else None

    class _Span:
        def __init__(self, eng, name):
            self.eng, self.name = eng, name

        def __enter__(self):
            if self.eng.nvtx:
                torch.cuda.nvtx.range_push("reid:" + self.name)
            if self.eng.timers is not None:
                self.a = torch.cuda.Event(enable_timing=True)
                self.b = torch.cuda.Event(enable_timing=True)
                self.a.record()
            return self

        def __exit__(self, *exc):
            if self.eng.nvtx:
                torch.cuda.nvtx.range_pop()
            if self.eng.timers is not None:
                self.b.record()
                self.eng.timers.setdefault(self.name, []).append((self.a, self.b))
            return False

    def span(self, name):
        return Engine._Span(self, name)

    def stage_ms(self):
        """{stage: total ms} since enable_timers(); synchronises."""
        torch.cuda.synchronize(self.device)
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in (self.timers or {}).items()}

    # ------------------------------------------------------------------ buffers
    def _empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def _zeros(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    # ------------------------------------------------------------------ K13
    def prepare(self, X: torch.Tensor, normalize: bool, X_aug: torch.Tensor | None = None) -> Features:
        """X: [n, d] float32 on the device (row-major).  X_aug (same shape): the second embedding pass of the callers'
        flip test-time augmentation; rows become (X + X_aug) / 2 inside the prologue kernel (reid_prepare_rows_tta)."""
        assert X.is_cuda and X.dtype == torch.float32 and X.dim() == 2 and X.stride(1) == 1
        if X_aug is not None:
            assert X_aug.is_cuda and X_aug.dtype == torch.float32 and X_aug.shape == X.shape and X_aug.stride(1) == 1
            n, d = X.shape
            dp = self.lib.reid_padded_width(d)
            x32, hi, lo = (self._empty((n, dp), t) for t in (torch.float32, torch.float16, torch.float16))
            sq, nrm = self._empty((n,), torch.float32), self._empty((n,), torch.float32)
            amax = (torch.full((1,), _ONE_BITS, dtype=torch.int32, device=self.device) if normalize
                    else self._empty((1,), torch.int32))
            _lib.call("reid_prepare_rows_tta", _ptr(X), _ptr(X_aug), n, d, X.stride(0), X_aug.stride(0),
                      1 if normalize else 0, _ptr(x32), _ptr(hi), _ptr(lo), dp, _ptr(sq), _ptr(nrm), _ptr(amax),
                      1 if normalize else 0, _stream())
            return Features(x32, hi, lo, sq, nrm, amax, d)
        n, d = X.shape
        dp = self.lib.reid_padded_width(d)
        x32 = self._empty((n, dp), torch.float32)
        hi = self._empty((n, dp), torch.float16)
        lo = self._empty((n, dp), torch.float16)
        sq = self._empty((n,), torch.float32)
        nrm = self._empty((n,), torch.float32)
        if normalize:
            # L2-normalised rows are bounded by 1: the split's power-of-two scale is fixed up front instead of taken from
            # the measured matrix-wide maximum, so a row's operands do not depend on which other rows it is prepared
            # with (the streamed evaluator prepares the rows chunk by chunk and must produce the same bits)
            amax = torch.full((1,), _ONE_BITS, dtype=torch.int32, device=self.device)
            _lib.call("reid_prepare_rows", _ptr(X), n, d, X.stride(0), 1, _ptr(x32), _ptr(hi), _ptr(lo), dp, _ptr(sq),
                      _ptr(nrm), _ptr(amax), 1, _stream())
        else:
            amax = self._empty((1,), torch.int32)
            _lib.call("reid_prepare_features", _ptr(X), n, d, X.stride(0), 0, _ptr(x32), _ptr(hi), _ptr(lo),
                      dp, _ptr(sq), _ptr(nrm), _ptr(amax), _stream())
        return Features(x32, hi, lo, sq, nrm, amax, d)

    # ------------------------------------------------------------------ K13 + K1, streamed from host memory
    @staticmethod
    def stream_bounds(N: int) -> list[int]:
        """Row boundaries (multiples of 256) of the chunks the streamed path works in.  The contraction of a chunk
        needs the rows before it, so its cost grows with the chunk's position; later chunks are smaller so that little
        work is left when the last byte has arrived."""
        if N < 4096:
            return [0, N]
        b = [0]
        for f in (0.40, 0.65, 0.80, 0.90):
            r = int(N * f) // 256 * 256
            if r > b[-1]:
                b.append(r)
        b.append(N)
        return b

    def prepare_dist_streamed(self, parts, normalize: bool = True):
        """parts: host float32 [n_i, d] tensors (pinned for a truly asynchronous copy), rows in order.  Copies them to
        the device chunk by chunk on a side stream and, as each chunk lands, prepares its rows and contracts the blocks
        of the symmetric distance matrix that have become computable (the diagonal block in upper-only form, and the
        rectangle above it with the mirrored statistics).  Same bits as prepare() + dist(upper_only=True) on the
        whole matrix.  Returns (D, rowmax_key, chunk_min, N)."""
        assert normalize and self.comm.world == 1 and self.engine != ENGINES["simt"]
        N = sum(int(p.shape[0]) for p in parts)
        d = int(parts[0].shape[1])
        dp = self.lib.reid_padded_width(d)
        X = self._empty((N, d), torch.float32)
        x32 = self._empty((N, dp), torch.float32)
        hi = self._empty((N, dp), torch.float16)
        lo = self._empty((N, dp), torch.float16)
        sq = self._empty((N,), torch.float32)
        nrm = self._empty((N,), torch.float32)
        amax = torch.full((1,), _ONE_BITS, dtype=torch.int32, device=self.device)
        ld = self.pitch(N)
        D = self._empty((N, ld), torch.float32)
        key = self._empty((N,), torch.int32)
        _lib.call("reid_rowmax_init", _ptr(key), N, _stream())
        cpitch = self.lib.reid_chunk_pitch(N)
        cmin = self._empty((N, cpitch), torch.float32)
        bounds = self.stream_bounds(N)
        cur = torch.cuda.current_stream(self.device)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        cs = self._copy_stream
        cs.wait_stream(cur)                       # X was allocated on the compute stream
        events = []
        with torch.cuda.stream(cs):
            r, ci = 0, 0
            for p in parts:
                assert p.dtype == torch.float32 and p.dim() == 2 and p.shape[1] == d
                off, n_p = 0, int(p.shape[0])
                while off < n_p:
                    take = min(n_p - off, bounds[ci + 1] - r)
                    X[r: r + take].copy_(p[off: off + take], non_blocking=True)
                    r += take
                    off += take
                    if r == bounds[ci + 1]:
                        ev = torch.cuda.Event()
                        ev.record(cs)
                        events.append(ev)
                        ci += 1
        st = _stream()
        for c in range(len(bounds) - 1):
            b0, b1 = bounds[c], bounds[c + 1]
            cur.wait_event(events[c])
            _lib.call("reid_prepare_rows", _ptr(X[b0:b1]), b1 - b0, d, X.stride(0), 1, _ptr(x32[b0:b1]), _ptr(hi[b0:b1]),
                      _ptr(lo[b0:b1]), dp, _ptr(sq[b0:b1]), _ptr(nrm[b0:b1]), _ptr(amax), 1, st)
            # diagonal block of this chunk: upper-only form inside the block
            _lib.call("reid_dist_block", _ptr(hi[b0:b1]), _ptr(lo[b0:b1]), hi.stride(0), _ptr(sq[b0:b1]), _ptr(amax),
                      b1 - b0, _ptr(hi[b0:b1]), _ptr(lo[b0:b1]), hi.stride(0), _ptr(sq[b0:b1]), _ptr(amax), b1 - b0,
                      dp, METRIC_SQEUCLIDEAN, _ptr(D[b0:b1, b0:]), ld, _ptr(key[b0:b1]), _ptr(cmin[b0:b1, b0 // 32:]),
                      cpitch, 2, None, None, st)
            if b0 > 0:
                # rows before the chunk x the chunk's columns: above the diagonal, mirrored statistics to the chunk's rows
                _lib.call("reid_dist_block", _ptr(hi[:b0]), _ptr(lo[:b0]), hi.stride(0), _ptr(sq[:b0]), _ptr(amax), b0,
                          _ptr(hi[b0:b1]), _ptr(lo[b0:b1]), hi.stride(0), _ptr(sq[b0:b1]), _ptr(amax), b1 - b0, dp,
                          METRIC_SQEUCLIDEAN, _ptr(D[:b0, b0:]), ld, _ptr(key), _ptr(cmin[:, b0 // 32:]), cpitch, 3,
                          _ptr(key[b0:b1]), _ptr(cmin[b0:b1]), st)
        return D, key, cmin[:, : (N + 31) // 32], N

    def evaluate_reranked_host(self, parts, q: int, pids: torch.Tensor, cams: torch.Tensor, k1: int, k2: int,
                               lam: float, max_rank: int = 50):
        """evaluate_reranked for features still in host memory (what evaluate.R1_mAP_eval.compute() holds): the copy
        overlaps the contraction.  One rank, normalised features, tensor-core engine."""
        with self.span("dist"):
            D, key, cmin, N = self.prepare_dist_streamed(parts, True)
        ids = (pids[:q].contiguous(), pids[q:].contiguous(), cams[:q].contiguous(), cams[q:].contiguous())
        _, res = self.rerank_from_panel(D, key, q, N, k1, k2, lam, None, chunk_min=cmin, score=(ids, max_rank),
                                        want_final=False, upper_only=True)
        return res

    # ------------------------------------------------------------------ K1/K2/K3
    @staticmethod
    def pitch(n: int) -> int:
        return (n + 31) // 32 * 32

    def dist(self, A: Features, B: Features, metric: int = METRIC_SQEUCLIDEAN, want_rowmax: bool = False,
             engine: int | None = None, want_chunk_min: bool = False, upper_only: bool = False,
             key_in: torch.Tensor | None = None):
        """Returns (D [m, pitch] float32 — use D[:, :n] —, rowmax_key [m] int32 or None)
        [, chunk_min [m, ceil(n/32)] float32 when want_chunk_min (tensor-core engine only)]."""
        assert A.dp == B.dp
        m, n = A.n, B.n
        ld = self.pitch(n)
        D = self._empty((m, ld), torch.float32)
        key = key_in                 # row maxima accumulate into a caller's array (panels of one row block): no init here
        if want_rowmax and key is None:
            key = self._empty((m,), torch.int32)
            _lib.call("reid_rowmax_init", _ptr(key), m, _stream())
        ra, rb = (A.sq, B.sq) if metric == METRIC_SQEUCLIDEAN else (A.nrm, B.nrm)
        engine = self.engine if engine is None else engine
        cmin = None
        if engine == ENGINES["simt"]:
            assert not want_chunk_min, "chunk minima come from the tensor-core epilogue"
            _lib.call("reid_dist_fp32", _ptr(A.x32), A.x32.stride(0), _ptr(ra), m, _ptr(B.x32), B.x32.stride(0),
                      _ptr(rb), n, A.dp, metric, _ptr(D), ld, _ptr(key), _stream())
        else:
            if want_chunk_min:
                cmin = self._empty((m, self.lib.reid_chunk_pitch(n)), torch.float32)[:, : (n + 31) // 32]
            sym_mode = (2 if upper_only else 1) if self.symmetric else 0      # per call: no process-wide kernel state
            _lib.call("reid_dist", _ptr(A.hi), _ptr(A.lo), A.hi.stride(0), _ptr(ra), _ptr(A.amax), m, _ptr(B.hi),
                      _ptr(B.lo), B.hi.stride(0), _ptr(rb), _ptr(B.amax), n, A.dp, metric, _ptr(D), ld, _ptr(key),
                      _ptr(cmin), sym_mode, _stream())
        return (D, key, cmin) if want_chunk_min else (D, key)

    def rowmax(self, D: torch.Tensor, cols: int) -> torch.Tensor:
        key = self._empty((D.shape[0],), torch.int32)
        _lib.call("reid_rowmax", _ptr(D), D.shape[0], cols, D.stride(0), _ptr(key), _stream())
        return key

    def rowmax_values(self, key: torch.Tensor) -> torch.Tensor:
        out = self._empty((key.shape[0],), torch.float32)
        _lib.call("reid_rowmax_decode", _ptr(key), key.shape[0], _ptr(out), _stream())
        return out

    def transposed(self, src: torch.Tensor, dst: torch.Tensor | None = None, accumulate: bool = False):
        """dst[i, j] (+)= src[j, i] for a square matrix; dst is pitched."""
        n = src.shape[0]
        assert src.shape[1] >= n and src.dtype == torch.float32
        if dst is None:
            dst = self._empty((n, self.pitch(n)), torch.float32)
        _lib.call("reid_transpose_add", _ptr(src), n, src.stride(0), _ptr(dst), dst.stride(0), int(accumulate),
                  _stream())
        return dst

    # ------------------------------------------------------------------ K5
    def topk_rows(self, D: torch.Tensor, cols: int, key: torch.Tensor, K: int, want_values=False, cap: int = 0,
                  chunk_min: torch.Tensor | None = None, upper_only: bool = False):
        rows = D.shape[0]
        nbr = self._empty((rows, K), torch.int32)
        val = self._empty((rows, K), torch.float32) if want_values else None
        if chunk_min is not None and upper_only and self.scatter_select:
            # scatter form: every stored element is offered to both of its rows, the unstored half is never gathered
            assert rows == cols
            nbytes = int(self.lib.reid_topk_scatter_workspace_bytes(rows, cap))
            ws = self._empty(((nbytes + 15) // 16 * 2,), torch.int64)
            _lib.call("reid_topk_rows_scatter", _ptr(D), rows, D.stride(0), _ptr(key), _ptr(chunk_min), K,
                      _ptr(nbr), _ptr(val), cap, _ptr(ws), ws.numel() * 8, _stream())
        elif chunk_min is not None:
            _lib.call("reid_topk_rows_hinted_tri", _ptr(D), rows, cols, D.stride(0), _ptr(key), _ptr(chunk_min), K,
                      _ptr(nbr), _ptr(val), cap, int(upper_only), _stream())
        elif cap:
            _lib.call("reid_topk_rows_capped", _ptr(D), rows, cols, D.stride(0), _ptr(key), K, _ptr(nbr), _ptr(val),
                      cap, _stream())
        else:
            _lib.call("reid_topk_rows", _ptr(D), rows, cols, D.stride(0), _ptr(key), K, _ptr(nbr), _ptr(val),
                      _stream())
        return (nbr, val) if want_values else nbr

    # ------------------------------------------------------------------ SURVEY §8 a13, first slice: streamed retrieval
    @staticmethod
    def column_blocks(g: int, cols_per_pass: int, K: int) -> list[tuple[int, int]]:
        """[c0, c1) gallery blocks of at most cols_per_pass columns (a tail shorter than K joins its predecessor:
        the selection kernels need at least K columns per block)."""
        starts = list(range(0, g, max(cols_per_pass, K)))
        if len(starts) > 1 and g - starts[-1] < K:
            starts.pop()
        return [(c0, starts[i + 1] if i + 1 < len(starts) else g) for i, c0 in enumerate(starts)]

    def retrieve_topk(self, Q: Features, G: Features, K: int, metric: int = METRIC_SQEUCLIDEAN,
                      rows_per_pass: int = 16384, cols_per_pass: int = 65536, on_panel=None,
                      rowmax: torch.Tensor | None = None):
        """Exact K nearest gallery rows of every query row, ascending in stable (distance, column) order, WITHOUT the
        q x g matrix ever existing (100k x 1M x 4 B = 400 GB): the contraction writes one [rows_per_pass, cols_per_pass]
        panel at a time together with its 32-column chunk minima, the hinted selection reads ~K chunks per row of it,
        and the per-block winners [rows, blocks * K] are selected once more.  Candidate position order equals global
        column order among equal distances, so the result is bit for bit what selection over the full matrix gives.
        Returns (idx int32 [q, K], dist float32 [q, K]); raw distances (the selection's divisor is 1).
        rowmax: optional int32 [q] (reid_rowmax_init'ed): receives every row's maximum over ALL gallery columns."""
        q, g = Q.n, G.n
        if not 1 <= K <= min(g, 256):
            raise ValueError(f"K={K} outside [1, min(g, 256)]")
        blocks = self.column_blocks(g, cols_per_pass, K)
        nb = len(blocks)
        unit = torch.full((min(q, rows_per_pass),), 0x3F800000, dtype=torch.int32, device=self.device)   # key of 1.0f
        out_i = self._empty((q, K), torch.int32)
        out_v = self._empty((q, K), torch.float32)
        sym = self.symmetric
        self.symmetric = False               # panels are rectangles of two different row sets
        try:
            for r0 in range(0, q, rows_per_pass):
                r1 = min(q, r0 + rows_per_pass)
                m = r1 - r0
                A = Q.rows(slice(r0, r1))
                if nb == 1:
                    cv, ci = out_v[r0:r1], out_i[r0:r1]
                else:
                    cv = self._empty((m, self.pitch(nb * K)), torch.float32)
                    ci = self._empty((m, nb * K), torch.int32)
                for b, (c0, c1) in enumerate(blocks):
                    with self.span("dist"):
                        D, _, cmin = self.dist(A, G.rows(slice(c0, c1)), metric, want_rowmax=False, want_chunk_min=True,
                                               key_in=None if rowmax is None else rowmax[r0:r1])
                    with self.span("select"):
                        idx, val = self.topk_rows(D, c1 - c0, unit[:m], K, want_values=True, chunk_min=cmin)
                        if nb == 1:
                            ci.copy_(idx)
                            cv.copy_(val)
                        else:
                            ci[:, b * K:(b + 1) * K] = idx + c0
                            cv[:, b * K:(b + 1) * K] = val
                    if on_panel is not None:
                        on_panel(D, r0, m, c0, c1)
                    del D, cmin
                if nb > 1:
                    with self.span("merge"):
                        pos, val = self.topk_rows(cv, nb * K, unit[:m], K, want_values=True)
                        out_i[r0:r1] = torch.gather(ci, 1, pos.long())
                        out_v[r0:r1] = val
        finally:
            self.symmetric = sym
        return out_i, out_v

    @staticmethod
    def sym_panel_bounds(N: int, rows_per_pass: int, Kw: int) -> list[int]:
        """Row-panel boundaries of the upper-half sweep: multiples of 256 (the contraction's block origins), a short tail
        (fewer than max(Kw, a quarter panel) rows) joined to the last panel -- its own diagonal block would give loose
        thresholds, and the selection needs Kw columns."""
        B = max(256, rows_per_pass // 256 * 256)
        starts = list(range(0, N, B))
        if len(starts) > 1 and N - starts[-1] < max(Kw, B // 4):
            starts.pop()
        return starts + [N]

    @staticmethod
    def sym_slab_rows(bounds: list[int], m: int, c0: int, c1: int, Kw: int) -> int:
        """Rows pushed at once into the columns [c0, c1) of a rectangle with m rows.  A column's threshold starts as the
        Kw-th distance inside its own diagonal block of t rows, so a panel of m rows holds ~Kw * m / t candidates for it;
        where that exceeds ~96 (a rectangle facing a short tail block) the rows go in slabs (multiples of 256) and the
        thresholds tighten in between.  Regular rectangles (t == m) are pushed whole at Kw = 64."""
        t_min = min(b1 - b0 for b0, b1 in zip(bounds, bounds[1:]) if b0 < c1 and b1 > c0)
        nslab = max(1, -(-(Kw * m) // (t_min * 96)))
        return (-(-m // nslab) + 255) // 256 * 256 if nslab > 1 else m

    def neighbour_lists_symmetric(self, F: Features, Kw: int, rows_per_pass: int = 16384, cols_per_pass: int = 65536):
        """The Kw nearest rows of EVERY row of F among all of F (raw squared distances, stable (distance, column)
        order) and every row's maximum, from the upper half of the symmetric N x N matrix only (csrc/stream_sym.cu):
        diagonal blocks in upper-only form (initial lists and thresholds), then the rectangles right of them, each
        serving its rows through the hinted selection and its columns through threshold-filtered pushes.  With P ranks
        the row panels are dealt out in snake order; every rank folds the pushes of ITS panels into its own copy of the
        lists and the copies are merged at the end.  Values are those of the materialised path's upper-only form.
        Returns (idx int32 [N, Kw], val float32 [N, Kw], rowmax_key int32 [N]) on every rank, or None when a column
        received more candidates than the buffers hold (the caller runs the full sweep instead)."""
        cm = self.comm
        r, P = cm.rank, cm.world
        N, dev, st = F.n, self.device, _stream
        CB = max(256, cols_per_pass // 256 * 256)
        bounds = self.sym_panel_bounds(N, rows_per_pass, Kw)
        starts = bounds[:-1]
        mine = [p for p in range(len(starts)) if _fabric.snake_owner(p, P) == r]
        key = self._empty((N,), torch.int32)
        _lib.call("reid_rowmax_init", _ptr(key), N, st())
        Mv = torch.full((N, Kw), float("inf"), dtype=torch.float32, device=dev)
        Mi = torch.full((N, Kw), 0x7FFFFFFF, dtype=torch.int32, device=dev)
        tau = torch.full((N,), float("inf"), dtype=torch.float32, device=dev)
        unit = torch.full((max(b1 - b0 for b0, b1 in zip(bounds, bounds[1:])),), _ONE_BITS, dtype=torch.int32, device=dev)
        hs = F.hi.stride(0)
        # ---- diagonal blocks: upper-only form, scatter selection -> the initial lists and thresholds of the own rows
        for p in mine:
            r0, r1 = bounds[p], bounds[p + 1]
            m = r1 - r0
            A = F.rows(slice(r0, r1))
            with self.span("dist"):
                D = self._empty((m, self.pitch(m)), torch.float32)
                cpitch = self.lib.reid_chunk_pitch(m)
                cmin = self._empty((m, cpitch), torch.float32)
                _lib.call("reid_dist_block", _ptr(A.hi), _ptr(A.lo), hs, _ptr(A.sq), _ptr(F.amax), m, _ptr(A.hi),
                          _ptr(A.lo), hs, _ptr(A.sq), _ptr(F.amax), m, F.dp, METRIC_SQEUCLIDEAN, _ptr(D), D.stride(0),
                          _ptr(key[r0:]), _ptr(cmin), cpitch, 2, None, None, st())
            with self.span("select"):
                idx, val = self.topk_rows(D, m, unit[:m], Kw, want_values=True, chunk_min=cmin[:, : (m + 31) // 32],
                                          upper_only=True)
                Mi[r0:r1] = idx + r0
                Mv[r0:r1] = val
                tau[r0:r1] = val[:, Kw - 1]
            del D, cmin
        if P > 1:
            cm.dist.all_reduce(tau, op=cm.dist.ReduceOp.MIN, group=cm.group)        # the other ranks hold +inf
        # ---- rectangles right of the diagonal blocks
        cap = min(self.sym_cap, 1024 - Kw)
        count = self._zeros((CB + Kw,), torch.int32)
        pbuf = self._empty(((CB + Kw) * cap,), torch.int64)
        cthr = self._empty(((CB + Kw + 31) // 32,), torch.int32)
        ovf = self._zeros((1,), torch.int32)
        for p in mine:
            r0, r1 = bounds[p], bounds[p + 1]
            m = r1 - r0
            A = F.rows(slice(r0, r1))
            blocks = [(r1 + c0, r1 + c1) for c0, c1 in self.column_blocks(N - r1, CB, Kw)] if r1 < N else []
            if not blocks:
                continue
            nb = len(blocks)
            cv = self._empty((m, self.pitch((nb + 1) * Kw)), torch.float32)
            ci = self._empty((m, (nb + 1) * Kw), torch.int32)
            for b, (c0, c1) in enumerate(blocks):
                n = c1 - c0
                Bf = F.rows(slice(c0, c1))
                with self.span("dist"):
                    D = self._empty((m, self.pitch(n)), torch.float32)
                    cpitch = self.lib.reid_chunk_pitch(n)
                    cmin = self._empty((m, cpitch), torch.float32)
                    _lib.call("reid_dist_block", _ptr(A.hi), _ptr(A.lo), hs, _ptr(A.sq), _ptr(F.amax), m, _ptr(Bf.hi),
                              _ptr(Bf.lo), hs, _ptr(Bf.sq), _ptr(F.amax), n, F.dp, METRIC_SQEUCLIDEAN, _ptr(D),
                              D.stride(0), _ptr(key[r0:]), _ptr(cmin), cpitch, 3, _ptr(key[c0:]), None, st())
                with self.span("select"):
                    idx, val = self.topk_rows(D, n, unit[:m], Kw, want_values=True, chunk_min=cmin[:, : (n + 31) // 32])
                    ci[:, (b + 1) * Kw:(b + 2) * Kw] = idx + c0
                    cv[:, (b + 1) * Kw:(b + 2) * Kw] = val
                with self.span("push"):
                    # slabs of rows: the thresholds tighten between slabs, which bounds the candidates a column can
                    # collect at once (rows much nearer to a column than its own block's -- queries against a gallery
                    # column -- would otherwise arrive by the hundred)
                    # expected candidates of a column per panel ~ Kw * m / (rows of the column's own diagonal block)
                    slab = self.sym_slab_rows(bounds, m, c0, c1, Kw)
                    for s0 in range(0, m, slab):
                        s1 = min(m, s0 + slab)
                        _lib.call("reid_sym_push", _ptr(D[s0:]), D.stride(0), s1 - s0, r0 + s0, n, c0, _ptr(cmin[s0:]),
                                  cpitch, _ptr(tau), _ptr(cthr), _ptr(count), _ptr(pbuf), cap, st())
                        _lib.call("reid_sym_fold", n, c0, Kw, _ptr(Mv), _ptr(Mi), _ptr(tau), _ptr(count), _ptr(pbuf),
                                  cap, _ptr(ovf), st())
                del D, cmin
            with self.span("merge"):
                # the rows' own lists so far (diagonal block + pushes of earlier panels: columns below r1) in front of the
                # rectangles' winners (columns from r1 on): position order == column order among equal distances
                ci[:, :Kw] = Mi[r0:r1]
                cv[:, :Kw] = Mv[r0:r1]
                pos, val = self.topk_rows(cv, (nb + 1) * Kw, unit[:m], Kw, want_values=True)
                Mi[r0:r1] = torch.gather(ci, 1, pos.long())
                Mv[r0:r1] = val
            del cv, ci
        if P > 1:
            with self.span("x_lists"):
                Fv, Fi = cm.gather_stack(Mv), cm.gather_stack(Mi)
                ovf_all = cm.gather_stack(ovf)
                cm.dist.all_reduce(key, op=cm.dist.ReduceOp.MAX, group=cm.group)
            self.last_sym_overflow = int(ovf_all.max().item())
            if self.last_sym_overflow:
                return None
            with self.span("merge"):
                _lib.call("reid_sym_merge", N, P, Kw, _ptr(Fv), _ptr(Fi), _ptr(Mv), _ptr(Mi), st())
            del Fv, Fi
        else:
            self.last_sym_overflow = int(ovf.item())
            if self.last_sym_overflow:
                return None
        return Mi, Mv, key

    def pair_distances(self, A: Features, B: Features, ri: torch.Tensor, ci: torch.Tensor,
                       metric: int = METRIC_SQEUCLIDEAN, chunk: int = 8192) -> torch.Tensor:
        """D[ri[p], ci[p]] for a list of (row of A, row of B) pairs with exactly the panels' arithmetic: the operand rows
        are gathered and contracted `chunk` x `chunk` on the tensor cores, the diagonal kept (an entry's value does not
        depend on the tile it is computed in).  ~1/chunk of each contraction is used, which is still cheap next to a
        panel sweep: 2 M pairs at d = 512 are 244 contractions of 0.15 ms."""
        npairs = int(ri.shape[0])
        out = self._empty((max(npairs, 1),), torch.float32)
        none = A.x32[:0]                 # the float32 rows are not operands of the tensor-core contraction
        simt = self.engine == ENGINES["simt"]
        sym = self.symmetric
        self.symmetric = False
        try:
            for a in range(0, npairs, chunk):
                b = min(npairs, a + chunk)
                ra, cb = ri[a:b], ci[a:b]
                fa = Features(A.x32[ra] if simt else none, A.hi[ra], A.lo[ra], A.sq[ra], A.nrm[ra], A.amax, A.d)
                fb = Features(B.x32[cb] if simt else none, B.hi[cb], B.lo[cb], B.sq[cb], B.nrm[cb], B.amax, B.d)
                Dm, _ = self.dist(fa, fb, metric)
                out[a:b] = torch.diagonal(Dm[:, : b - a])
                del Dm, fa, fb
        finally:
            self.symmetric = sym
        return out[:npairs]

    def evaluate_streamed(self, Q: Features, G: Features, q_pids, g_pids, q_cams, g_cams, max_rank: int = 50,
                          K: int = 0, metric: int = METRIC_SQEUCLIDEAN, rows_per_pass: int = 16384,
                          cols_per_pass: int = 65536, q_total: int | None = None):
        """evaluate.py:40-88 (CMC / mAP with same-id same-camera junk removed) for a q x g problem too large to hold:
        the distances of every query's same-identity columns are contracted as gathered pairs first (pair_distances) and
        sorted into match keys; then the contraction is swept ONCE over [rows_per_pass x cols_per_pass] panels
        (stream_score.cu), each panel feeding the exact top-K selection of retrieve_topk (K > 0) and the count of the
        entries in front of every match.  Q / q_pids / q_cams are this rank's queries.
        Returns ((cmc, mAP, n_valid, ap, first_hit), (idx, dist) or None)."""
        nq, g = Q.n, G.n
        keys, start, lst, T = self.pid_index(g_pids)
        glo = self._empty((nq,), torch.int32)
        glen = self._empty((nq,), torch.int32)
        _lib.call("reid_stream_group_lookup", _ptr(q_pids), nq, _ptr(keys), _ptr(start), T, _ptr(glo), _ptr(glen), _stream())
        goff = self._zeros((nq + 1,), torch.int64)
        goff[1:] = torch.cumsum(glen, 0, dtype=torch.int64)
        # the evaluation's one host read before the sweep: buffer size, and the largest identity group -- refused here,
        # not after a full contraction sweep
        sizes = torch.stack([goff[-1], glen.max().to(torch.int64) if nq else goff[-1]]).tolist()
        total, longest = int(sizes[0]), int(sizes[1])
        if longest > GROUP_MAX:
            raise _lib.ReidError(f"an identity has {longest} gallery samples; the streamed scorer holds at most "
                                 f"{GROUP_MAX} per query")
        mkeys = self._empty((max(total, 1),), torch.int64)
        bins = self._zeros((max(total, 1),), torch.int32)
        nmatch = self._empty((nq,), torch.int32)
        overflow = self._zeros((1,), torch.int32)
        # The distances of every query's same-identity columns are contracted directly (gathered operand rows, the
        # panels' arithmetic), so the sorted match keys exist BEFORE the sweep and ONE sweep over the panels does the
        # top-K selection and the counting (the first version swept twice: gather, then count).
        with self.span("pairs"):
            if total:
                qi = torch.repeat_interleave(torch.arange(nq, device=self.device), glen.long(), output_size=total)
                pos = torch.arange(total, device=self.device) - goff[:-1][qi]
                gd = self.pair_distances(Q, G, qi, lst[glo.long()[qi] + pos].long(), metric)
            else:
                gd = self._empty((1,), torch.float32)
            _lib.call("reid_stream_group_sort", nq, _ptr(glo), _ptr(glen), _ptr(goff), _ptr(lst), _ptr(gd), _ptr(q_cams),
                      _ptr(g_cams), _ptr(mkeys), _ptr(nmatch), _ptr(overflow), longest, _stream())

        def count(D, r0, m, c0, c1):
            _lib.call("reid_stream_count_panel", _ptr(D), D.stride(0), m, c0, c1, _ptr(goff[r0:]), _ptr(mkeys),
                      _ptr(nmatch[r0:]), _ptr(bins), longest, _stream())

        top = None
        with self.span("sweep"):
            if K > 0:
                top = self.retrieve_topk(Q, G, K, metric, rows_per_pass, cols_per_pass, on_panel=count)
            else:
                self._sweep(Q, G, metric, rows_per_pass, cols_per_pass, count)

        def launch(ap, first, hist, nvalid):
            _lib.call("reid_stream_score", nq, _ptr(glen), _ptr(goff), _ptr(mkeys), _ptr(nmatch), _ptr(bins), max_rank,
                      _ptr(ap), _ptr(first), _ptr(hist), _ptr(nvalid), _stream())
        res = self._score_common(launch, nq, max_rank, q_total, span="score")
        big = int(overflow.item())
        if big:
            raise _lib.ReidError(f"an identity has {big} gallery samples; the streamed scorer holds at most "
                                 f"{GROUP_MAX} per query")
        return res, top

    def _sweep(self, Q: Features, G: Features, metric: int, rows_per_pass: int, cols_per_pass: int, on_panel):
        """One pass of the contraction over [rows_per_pass x cols_per_pass] panels of Q x G; on_panel(D, r0, m, c0, c1)."""
        sym = self.symmetric
        self.symmetric = False
        try:
            for r0 in range(0, Q.n, rows_per_pass):
                r1 = min(Q.n, r0 + rows_per_pass)
                A = Q.rows(slice(r0, r1))
                for c0, c1 in self.column_blocks(G.n, cols_per_pass, 1):
                    with self.span("dist"):
                        D, _ = self.dist(A, G.rows(slice(c0, c1)), metric)
                    with self.span("panel"):
                        on_panel(D, r0, r1 - r0, c0, c1)
                    del D
        finally:
            self.symmetric = sym

    # ------------------------------------------------------------------ re-ranking with no dense matrix
    def evaluate_streamed_reranked(self, allf: Features, q: int, pids: torch.Tensor, cams: torch.Tensor, k1: int,
                                   k2: int, lam: float, topk: int = 100, max_rank: int = 50, Kw: int = 64,
                                   Kd: int | None = None, cap_o: int = 32768, rows_per_pass: int = 16384,
                                   cols_per_pass: int = 65536, detail: dict | None = None):
        """reranking.py:29-100 + evaluate.py:40-88 for a problem whose N x N (and q x g) matrix cannot exist (BASELINE
        config 5: 100k queries x 1M gallery): k-reciprocal re-ranked CMC / mAP and the exact top-`topk` of every
        re-ranked row.  allf: the prepared N = q + g rows (queries first), pids / cams [N] int64 on the device.

          A  N x N panel sweep (this rank's rows, cyclic): per row the Kw nearest by raw distance and the row maximum
             -> all-gather -> lists re-keyed by D / max: initial_rank[:, :K]                          (reranking.py:46-48)
          B  reciprocal sets, V weights (distances looked up in the wide lists), k2 expansion, inverted index -- the
             sparse kernels of the materialised path, replicated on every rank                      (reranking.py:51-82)
          C  own queries: ordered float16 Jaccard chain into per-query overlap lists                (reranking.py:84-93)
          D  distances of every query's same-identity columns, contracted as gathered pairs (pair_distances)
          E  final keys of those entries, raw-distance thresholds of every match                        (reranking.py:95)
          F  ONE q x g sweep: top-Kd by raw distance, exact distances of the overlap columns, every entry binned as if
             untouched;  G  corrections by the special entries, AP / CMC; exact top-k merge
        Same arithmetic as the materialised path with the full (non-symmetric) tile schedule: identical results whenever
        `diag["v_misses"]` is 0 (tests/test_gpu_retrieval.py).  Returns ((cmc, mAP, n_valid, ap, first_hit),
        (idx int32 [nq_local, topk], final float32 [nq_local, topk]), diag)."""
        cm = self.comm
        r, P = cm.rank, cm.world
        N = allf.n
        g = N - q
        K = max(k1 + 1, k2)
        Kw = min(max(Kw, K), N, 256)
        Kd = min(g, 256, topk + 28) if Kd is None else min(Kd, g, 256)
        if not (1 <= topk <= Kd and N >= K):
            raise ValueError(f"topk={topk} / K={K} do not fit q={q}, g={g}")
        st = _stream
        # ---- A
        lists = None
        self.last_sym_overflow = 0
        with self.span("nn_sweep"):
            if self.sym_sweep and self.symmetric and self.engine != ENGINES["simt"] and P * Kw <= 1024:
                # half the matrix: values are the upper-triangle ones, as in the materialised path's upper-only form
                lists = self.neighbour_lists_symmetric(allf, Kw, rows_per_pass, cols_per_pass)
            if lists is None:
                A = allf.rows(slice(r, N, P))
                key_l = self._empty((A.n,), torch.int32)
                _lib.call("reid_rowmax_init", _ptr(key_l), A.n, st())
                widx_l, wval_l = self.retrieve_topk(A, allf, Kw, METRIC_SQEUCLIDEAN, rows_per_pass, cols_per_pass,
                                                    rowmax=key_l)
        if lists is not None:
            widx, wval, key = lists
        else:
            with self.span("x_lists"):
                widx, wval, key = cm.gather_rows(widx_l, N), cm.gather_rows(wval_l, N), cm.gather_rows(key_l, N)
        upper = lists is not None
        # ---- B (replicated)
        flags = self._zeros((5,), torch.int32)      # rekey ambiguous | V misses | overlap overflow | group overflow | topk ambiguous
        with self.span("krecip"):
            nbr = self._empty((N, K), torch.int32)
            _lib.call("reid_lists_rekey", _ptr(widx), _ptr(wval), N, Kw, _ptr(key), K, _ptr(nbr), _ptr(flags[0:]), st())
            rfull, nfull, rhalf, nhalf = self.krecip_sets(nbr, k1)
            cap1 = v_row_capacity(k1)
            Vc = self._empty((N, cap1), torch.int32)
            Vv = self._empty((N, cap1), torch.float16)
            Vn = self._empty((N,), torch.int32)
            miss_cap = max(1 << 16, N * 2)
            miss_out = self._empty((miss_cap,), torch.int64)

            def v_pass(mt_keys, mt_vals, mt_n, out, cap):
                flags[1:2].zero_()
                _lib.call("reid_krecip_V_lists", _ptr(rfull), _ptr(nfull), _ptr(rhalf), _ptr(nhalf), N, k1, _ptr(key),
                          _ptr(widx), _ptr(wval), Kw, _ptr(allf.x32), allf.x32.stride(0), _ptr(allf.sq), _ptr(Vc),
                          _ptr(Vv), _ptr(Vn), _ptr(flags[1:]), _ptr(mt_keys), _ptr(mt_vals), mt_n, _ptr(out), cap, st())

            v_pass(None, None, 0, miss_out, miss_cap)
            n_miss = int(flags[1].item())                  # pairs of E_i outside row i's wide list (host read)
            pairs_recomputed = n_miss
            if n_miss > miss_cap:
                raise _lib.ReidError(f"streamed re-ranking: {n_miss} V pairs fell outside the wide lists (Kw={Kw}); raise Kw")
            if n_miss:
                # exactly those pairs on the tensor cores: gathered operand rows, 2048 pairs per contraction, diagonals
                mt_keys = torch.sort(miss_out[:n_miss]).values
                mt_vals = self._zeros((n_miss,), torch.float32)
                ri, ci = (mt_keys >> 32), (mt_keys & 0xFFFFFFFF)
                if upper:        # the value row i holds for column e: D[i, e] from i's 256-wide diagonal block on, else D[e, i]
                    swap = ci < (ri & ~255)
                    ri, ci = torch.where(swap, ci, ri), torch.where(swap, ri, ci)
                sym = self.symmetric
                self.symmetric = False
                try:
                    chunk = 8192                 # pairs per contraction: [8192 x 8192] = 1 024 tiles, 8 192 values kept
                    starts = list(range(0, n_miss, chunk))
                    for a in starts[r::P]:       # ranks take the chunks in turn; the values are summed across ranks below
                        b = min(n_miss, a + chunk)
                        none = allf.x32[:0]      # the float32 rows are not operands of the tensor-core contraction
                        fa = Features(none, allf.hi[ri[a:b]], allf.lo[ri[a:b]], allf.sq[ri[a:b]], allf.nrm[ri[a:b]],
                                      allf.amax, allf.d)
                        fb = Features(none, allf.hi[ci[a:b]], allf.lo[ci[a:b]], allf.sq[ci[a:b]], allf.nrm[ci[a:b]],
                                      allf.amax, allf.d)
                        Dm, _ = self.dist(fa, fb, METRIC_SQEUCLIDEAN)
                        mt_vals[a:b] = torch.diagonal(Dm[:, : b - a])
                        del Dm, fa, fb
                finally:
                    self.symmetric = sym
                cm.all_reduce_sum(mt_vals)       # every value was written by one rank, the others hold exact zeros
                v_pass(mt_keys, mt_vals, n_miss, None, 0)
        if k2 != 1:
            with self.span("qe"):
                Qc, Qv, qoff, Qn, nnz_bound = self.query_expansion(nbr, k2, Vc, Vv, Vn, N, 0, 1)
        else:
            qoff = self._empty((N + 1,), torch.int64)
            _lib.call("reid_ell_offsets", N, cap1, _ptr(qoff), st())
            Qc, Qv, Qn = Vc.reshape(-1), Vv.reshape(-1), Vn
            nnz_bound = N * cap1
        with self.span("csc"):
            colptr, csc_ent = self.inverted_index(Qc, Qv, qoff, Qn, N, nnz_bound)
        # ---- C
        nq = cm.local_count(q)
        qsel = slice(r, q, P)
        q_pids, g_pids = pids[:q][qsel].contiguous(), pids[q:].contiguous()
        q_cams, g_cams = cams[:q][qsel].contiguous(), cams[q:].contiguous()
        with self.span("overlap"):
            ocol = self._empty((max(nq, 1) * cap_o,), torch.int32)
            ot = self._empty((max(nq, 1) * cap_o,), torch.float16)
            od = self._zeros((max(nq, 1) * cap_o,), torch.float32)
            ocnt = self._empty((max(nq, 1),), torch.int32)
            # column ranges: sized by the largest overlap list the previous call saw (x2 head room inside a range's 8 192
            # slots); an overflow with hinted ranges is retried once with the default before it is reported
            hint = self._overlap_hint.get((N, q, k1, k2), 0)
            nrange = max(1, -(-2 * hint // 8192)) if hint else 0
            for attempt in (nrange, 0):
                flags[2:3].zero_()
                _lib.call("reid_overlap_lists", _ptr(Qc), _ptr(Qv), _ptr(qoff), _ptr(Qn), _ptr(colptr), _ptr(csc_ent),
                          r, P, nq, q, N, cap_o, attempt, _ptr(ocol), _ptr(ot), _ptr(ocnt), _ptr(flags[2:]), st())
                if attempt == 0 or int(flags[2].item()) == 0:
                    break
        # ---- D
        Qf, Gf = allf.rows(qsel), allf.rows(slice(q, N))
        keys, start, lst, T = self.pid_index(g_pids)
        glo = self._empty((max(nq, 1),), torch.int32)
        glen = self._empty((max(nq, 1),), torch.int32)
        _lib.call("reid_stream_group_lookup", _ptr(q_pids), nq, _ptr(keys), _ptr(start), T, _ptr(glo), _ptr(glen), st())
        goff = self._zeros((nq + 1,), torch.int64)
        goff[1:] = torch.cumsum(glen[:nq], 0, dtype=torch.int64)
        sizes = torch.stack([goff[-1], glen[:nq].max().to(torch.int64) if nq else goff[-1]]).tolist()   # one host read
        total, longest = int(sizes[0]), int(sizes[1])
        if longest > GROUP_MAX:
            raise _lib.ReidError(f"an identity has {longest} gallery samples; the streamed scorer holds at most "
                                 f"{GROUP_MAX} per query")
        # the distances of the same-identity columns are contracted as gathered pairs (the panels' bits), so that the match
        # thresholds exist before the q x g sweep and ONE sweep serves selection, the overlap columns and the counting
        with self.span("pairs"):
            if total:
                qi = torch.repeat_interleave(torch.arange(nq, device=self.device), glen[:nq].long(), output_size=total)
                pos = torch.arange(total, device=self.device) - goff[:-1][qi]
                gd = self.pair_distances(Qf, Gf, qi, lst[glo[:nq].long()[qi] + pos].long())
            else:
                gd = self._empty((1,), torch.float32)
        # ---- E
        c1 = float(np.float16(1 - lam))
        c2 = float(np.float32(lam))
        gkey = self._empty((max(total, 1),), torch.int64)
        gjunk = self._empty((max(total, 1),), torch.uint8)
        mk = self._empty((max(total, 1),), torch.int64)
        thrA = self._empty((max(total, 1),), torch.float32)
        thrB = self._empty((max(total, 1),), torch.float32)
        nmatch = self._empty((max(nq, 1),), torch.int32)
        bins = self._zeros((max(total, 1),), torch.int32)
        with self.span("rr_prepare"):
            _lib.call("reid_rr_prepare", nq, r, P, _ptr(glo), _ptr(glen), _ptr(goff), _ptr(lst), _ptr(gd), _ptr(q_cams),
                      _ptr(g_cams), _ptr(key), c1, c2, cap_o, _ptr(ocol), _ptr(ot), _ptr(ocnt), _ptr(gkey), _ptr(gjunk),
                      _ptr(mk), _ptr(thrA), _ptr(thrB), _ptr(nmatch), _ptr(flags[3:]), longest, st())

        # ---- F: the q x g sweep
        def on_panel(D, r0, m, c0, c1):
            _lib.call("reid_stream_list_gather", _ptr(D), D.stride(0), m, c0, c1, cap_o, _ptr(ocol[r0 * cap_o:]),
                      _ptr(ocnt[r0:]), _ptr(od[r0 * cap_o:]), st())
            _lib.call("reid_stream_count_panel_rr", _ptr(D), D.stride(0), m, c0, c1, _ptr(goff[r0:]), _ptr(mk), _ptr(thrA),
                      _ptr(thrB), _ptr(nmatch[r0:]), _ptr(bins), longest, st())

        with self.span("sweep"):
            tidx, tval = self.retrieve_topk(Qf, Gf, Kd, METRIC_SQEUCLIDEAN, rows_per_pass, cols_per_pass, on_panel=on_panel)

        # ---- G
        def launch(ap, first, hist, nvalid):
            _lib.call("reid_stream_score_rr", nq, r, P, _ptr(glo), _ptr(glen), _ptr(goff), _ptr(lst), _ptr(gd), _ptr(gkey),
                      _ptr(gjunk), _ptr(mk), _ptr(thrA), _ptr(thrB), _ptr(nmatch), _ptr(bins), cap_o, _ptr(ocol), _ptr(ot),
                      _ptr(od), _ptr(ocnt), _ptr(q_pids), _ptr(g_pids), _ptr(key), c1, c2, max_rank, _ptr(ap), _ptr(first),
                      _ptr(hist), _ptr(nvalid), longest, st())
        res = self._score_common(launch, nq, max_rank, q, span="score")
        with self.span("rr_topk"):
            out_i = self._empty((max(nq, 1), topk), torch.int32)
            out_v = self._empty((max(nq, 1), topk), torch.float32)
            _lib.call("reid_rr_topk", nq, r, P, _ptr(tidx), _ptr(tval), Kd, g, cap_o, _ptr(ocol), _ptr(ot), _ptr(od),
                      _ptr(ocnt), _ptr(key), c1, c2, topk, _ptr(out_i), _ptr(out_v), _ptr(flags[4:]), st())
        f = [int(v) for v in flags.tolist()]
        osz = (torch.stack([ocnt[:nq].max(), ocnt[:nq].sum()]).tolist() if nq else [0, 0])
        self._overlap_hint[(N, q, k1, k2)] = cm.max_int(int(osz[0]), self.device)     # same ranges on every rank next time
        diag = {"overlap_targets_max": int(osz[0]), "overlap_targets_mean": (osz[1] / nq if nq else 0.0),
                "rekey_ambiguous": f[0], "v_misses": f[1], "v_pairs_recontracted": pairs_recomputed,
                "overlap_overflow": f[2], "group_overflow": f[3], "topk_ambiguous": f[4], "Kw": Kw, "Kd": Kd,
                "nn_sweep": "upper half" if upper else "full", "sym_overflow": self.last_sym_overflow}
        if f[2] or f[3]:
            raise _lib.ReidError(f"streamed re-ranking: {f[2]} queries touch more than cap_o={cap_o} gallery columns, "
                                 f"largest identity group {f[3]} (limit {GROUP_MAX})")
        if f[0] or f[4]:
            raise _lib.ReidError(f"streamed re-ranking: candidate lists too short for exactness ({f[0]} rows at Kw={Kw}, "
                                 f"{f[4]} queries at Kd={Kd}); raise Kw / Kd")
        if detail is not None:
            detail.update(nbr=nbr, Vc=Vc, Vv=Vv, Vn=Vn, key=key, ocol=ocol, ot=ot, od=od, ocnt=ocnt, tidx=tidx, tval=tval)
        return res, (out_i[:nq], out_v[:nq]), diag

    # ------------------------------------------------------------------ K6
    def krecip_sets(self, nbr: torch.Tensor, k1: int):
        N, K = nbr.shape
        h = half_k(k1)
        rfull = self._empty((N, k1 + 1), torch.int32)
        nfull = self._empty((N,), torch.int32)
        rhalf = self._empty((N, h), torch.int32)
        nhalf = self._empty((N,), torch.int32)
        _lib.call("reid_krecip_sets", _ptr(nbr), N, K, k1, _ptr(rfull), _ptr(nfull), _ptr(rhalf), _ptr(nhalf),
                  _stream())
        return rfull, nfull, rhalf, nhalf

    def krecip_V(self, sets, k1: int, D: torch.Tensor, key: torch.Tensor, row0=0, row_stride=1, upper_only=False):
        rfull, nfull, rhalf, nhalf = sets
        rows = D.shape[0]
        cap1 = v_row_capacity(k1)
        Vc = self._empty((rows, cap1), torch.int32)
        Vv = self._empty((rows, cap1), torch.float16)
        Vn = self._empty((rows,), torch.int32)
        if upper_only:
            assert row0 == 0 and row_stride == 1
            _lib.call("reid_krecip_V_tri", _ptr(rfull), _ptr(nfull), _ptr(rhalf), _ptr(nhalf), rows, k1,
                      _ptr(D), D.stride(0), _ptr(key), _ptr(Vc), _ptr(Vv), _ptr(Vn), 1, _stream())
        else:
            _lib.call("reid_krecip_V", _ptr(rfull), _ptr(nfull), _ptr(rhalf), _ptr(nhalf), row0, row_stride, rows, k1,
                      _ptr(D), D.stride(0), _ptr(key), _ptr(Vc), _ptr(Vv), _ptr(Vn), _stream())
        return Vc, Vv, Vn

    def _qe_cap_hint(self) -> int:
        """Power-of-two merge-buffer size covering the longest union seen by the previous no-sync call (256 at first)."""
        if not self._capturing and self._ubmax_pending and self._ubmax_event.query():
            self._ubmax_seen = int(self._ubmax_pinned[0])
            self._ubmax_pending = False
        cap = 32
        while cap < min(self._ubmax_seen, 4096):
            cap *= 2
        return cap

    # ------------------------------------------------------------------ K7
    def query_expansion(self, nbr, k2, Vc, Vv, Vn, rows, row0=0, row_stride=1, cap_sort: int = 0, extra=None):
        """Returns (Qc, Qv, qoff[rows+1] int64, Qn[rows], pool size) for the row window.  `extra`: a one-element
        device tensor read back with the sizes (same D2H); its value is left in self.last_extra."""
        K = nbr.shape[1]
        cap1 = Vc.shape[1]
        ub = self._empty((rows,), torch.int32)
        _lib.call("reid_qe_bound", _ptr(nbr), K, k2, _ptr(Vn), row0, row_stride, rows, _ptr(ub), _stream())
        qoff = self._empty((rows + 1,), torch.int64)
        _lib.call("reid_exclusive_scan", _ptr(ub), rows, _ptr(qoff), _stream())
        # 16-byte D2H: pool size and the longest union (sizes the per-warp sort buffers)
        bound = rows * k2 * cap1                      # structural: a union of k2 rows of at most cap1 entries each
        if extra is None and not cap_sort and 0 < bound * 14 <= self.qe_nosync_bytes:
            # No host read in the middle of the step: the pools are sized by the structural bound (6 B per entry here,
            # 8 B in the inverted index) and the per-warp merge buffers by the longest union of the PREVIOUS call, which
            # arrives through a pinned buffer; rows longer than that take the kernel's exact k2-way merge fallback.
            total = bound
            cap_sort = self._qe_cap_hint()
            if not self._capturing:               # (the hint's host round trip stays outside a captured step)
                self._ubmax_pinned.copy_(ub.max().to(torch.int64).reshape(1), non_blocking=True)
                self._ubmax_event.record()
                self._ubmax_pending = True
        else:
            stats = [qoff[rows].reshape(1), ub.max().to(torch.int64).reshape(1)] if rows else []
            if extra is not None:
                stats.append(extra.to(torch.int64).reshape(1))
            vals = [int(v) for v in torch.cat(stats).tolist()] if stats else []
            total, ub_max = (vals[0], vals[1]) if rows else (0, 0)
            self.last_extra = vals[-1] if extra is not None else None
            if not cap_sort:
                cap_sort = 32
                while cap_sort < min(ub_max, 4096):
                    cap_sort *= 2
        Qc = self._empty((max(total, 1),), torch.int32)
        Qv = self._empty((max(total, 1),), torch.float16)
        Qn = self._empty((rows,), torch.int32)
        _lib.call("reid_qe_k2_capped", _ptr(nbr), K, k2, _ptr(Vc), _ptr(Vv), _ptr(Vn), cap1, row0, row_stride,
                  rows, _ptr(qoff), cap_sort, _ptr(Qc), _ptr(Qv), _ptr(Qn), _stream())
        return Qc, Qv, qoff, Qn, total

    def _query_expansion_sharded(self, nbr, k2, Vc, Vv, Vn, Vn_true, N: int):
        """K7 for this rank's rows + the exchange.  Every rank bounds the union size of EVERY row (cheap), so pool
        sizes and offsets of all ranks are known everywhere without a collective: one 8(P+1)-byte D2H sizes the
        buffers, two all-gathers move the pools (int32: columns | counts, float16: values; same pitch, so one offset
        array addresses both)."""
        cm = self.comm
        r, P = cm.rank, cm.world
        nloc, nm = cm.local_count(N), cm.nmax(N)
        K, cap1 = nbr.shape[1], Vc.shape[1]
        with self.span("qe"):
            ub = self._zeros((nm * P,), torch.int32)
            _lib.call("reid_qe_bound", _ptr(nbr), K, k2, _ptr(Vn), 0, 1, N, _ptr(ub), _stream())
            totals, excl = cm.pool_layout(ub, P)
            stats = torch.cat([totals, ub.max().to(torch.int64).reshape(1), Vn_true.max().to(torch.int64).reshape(1)]).tolist()
            tmax, total, ub_max, widest = max(stats[:P]), stats[r], stats[P], stats[P + 1]
            pitch = (tmax + nm + 7) // 8 * 8                         # pool | per-row counts, in elements
            ibuf = self._zeros((pitch,), torch.int32)
            hbuf = self._zeros((pitch,), torch.float16)
            qoff_l = torch.cat([excl[r, :nloc], totals[r: r + 1]]).contiguous()
            Qn_l = ibuf[tmax: tmax + nloc]
            cap_sort = 32
            while cap_sort < min(ub_max, 4096):
                cap_sort *= 2
            _lib.call("reid_qe_k2_capped", _ptr(nbr), K, k2, _ptr(Vc), _ptr(Vv), _ptr(Vn), cap1, r, P, nloc,
                      _ptr(qoff_l), cap_sort, _ptr(ibuf), _ptr(hbuf), _ptr(Qn_l), _stream())
        with self.span("x_qe"):
            iall = cm.gather_stack(ibuf)                             # [P, pitch]
            hall = cm.gather_stack(hbuf)
            Qn = cm.interleave(iall[:, tmax: tmax + nm], N).contiguous()
            base = torch.arange(P, device=self.device, dtype=torch.int64).unsqueeze(1) * pitch
            qoff = torch.cat([cm.interleave(excl + base, N), excl.new_zeros(1)]).contiguous()
        return iall.reshape(-1), hall.reshape(-1), qoff, Qn, pitch * P, widest

    # ------------------------------------------------------------------ K8
    def inverted_index(self, Qc, Qv, qoff, Qn, N: int, nnz_bound: int):
        colcnt = self._zeros((N,), torch.int32)
        _lib.call("reid_csc_count", _ptr(Qc), _ptr(qoff), _ptr(Qn), N, _ptr(colcnt), _stream())
        colptr = self._empty((N + 1,), torch.int64)
        _lib.call("reid_exclusive_scan", _ptr(colcnt), N, _ptr(colptr), _stream())
        cursor = self._zeros((N,), torch.int32)
        csc_ent = self._empty((max(nnz_bound, 1),), torch.int64)     # row | float16 bits << 32
        _lib.call("reid_csc_fill", _ptr(Qc), _ptr(Qv), _ptr(qoff), _ptr(Qn), N, _ptr(colptr), _ptr(cursor),
                  _ptr(csc_ent), _stream())
        return colptr, csc_ent

    def _jaccard_ws(self, nq_local: int, N: int, force_dense: bool = False):
        if force_dense or not self.sparse_cap:
            return None, 0
        nbytes = int(self.lib.reid_jaccard_ws_bytes(nq_local, N, self.sparse_cap))
        if nbytes <= 0:
            return None, 0
        return self._empty((nbytes,), torch.uint8), nbytes

    # ------------------------------------------------------------------ K9 + K10
    def jaccard_mix(self, Qc, Qv, qoff, Qn, csc, D, key, q: int, N: int, lam: float, nq_local: int, q0=0,
                    q_stride=1, want_jaccard=False):
        colptr, csc_ent = csc
        g = N - q
        ldf = self.pitch(g)
        out = self._empty((nq_local, ldf), torch.float32)
        J = self._empty((nq_local, N), torch.float16) if want_jaccard else None
        c1 = float(np.float16(1 - lam))            # fp16 array * python float: scalar cast to fp16 first
        c2 = float(np.float32(lam))                # fp32 array * python float
        ws, ws_bytes = self._jaccard_ws(nq_local, N, force_dense=want_jaccard)
        _lib.call("reid_jaccard_mix", _ptr(Qc), _ptr(Qv), _ptr(qoff), _ptr(Qn), _ptr(colptr), _ptr(csc_ent),
                  _ptr(D), D.stride(0), _ptr(key), q0, q_stride, nq_local, q, N, c1, c2, _ptr(out),
                  ldf, _ptr(J), N, _ptr(ws), ws_bytes, self.sparse_cap, _stream())
        return (out, J) if want_jaccard else out

    # ------------------------------------------------------------------ identity index + fused K9–K12
    def pid_index(self, g_pids: torch.Tensor):
        """Groups the gallery columns by pid (evaluate.py:43, :53-55) for reid_jaccard_score."""
        g = g_pids.shape[0]
        T = self.lib.reid_pid_index_slots(g)
        keys = self._empty((T,), torch.int64)
        cnt = self._empty((T + 1,), torch.int32)
        cursor = self._empty((T + 1,), torch.int32)
        start = self._empty((T + 2,), torch.int64)
        lst = self._empty((max(g, 1),), torch.int32)
        _lib.call("reid_pid_index", _ptr(g_pids), g, T, _ptr(keys), _ptr(cnt), _ptr(start), _ptr(cursor), _ptr(lst),
                  _stream())
        return keys, start, lst, T

    def jaccard_score(self, Qc, Qv, qoff, Qn, csc, D, key, q: int, N: int, lam: float, nq_local: int, q0, q_stride,
                      ids, max_rank: int, want_final: bool, fabric=None):
        """Fused K9–K12 for this rank's queries.  ids = (q_pids_local, g_pids, q_cams_local, g_cams).
        Returns (final or None, (cmc, mAP, n_valid, ap_all, first_all))."""
        colptr, csc_ent = csc
        q_pids, g_pids, q_cams, g_cams = ids
        g = N - q
        out, ldf = None, 0
        if want_final:
            ldf = self.pitch(g)
            out = self._empty((nq_local, ldf), torch.float32)
        keys, start, lst, T = self.pid_index(g_pids)
        c1 = float(np.float16(1 - lam))
        c2 = float(np.float32(lam))
        ws, ws_bytes = self._jaccard_ws(nq_local, N)

        def launch(ap, first, hist, nvalid):
            _lib.call("reid_jaccard_score", _ptr(Qc), _ptr(Qv), _ptr(qoff), _ptr(Qn), _ptr(colptr), _ptr(csc_ent),
                      _ptr(D), D.stride(0), _ptr(key), q0, q_stride, nq_local, q, N, c1, c2, _ptr(out),
                      ldf, _ptr(q_pids), _ptr(q_cams), _ptr(g_cams), _ptr(keys), _ptr(start), _ptr(lst), T, max_rank,
                      _ptr(ap), _ptr(first), _ptr(hist), _ptr(nvalid), _ptr(ws), ws_bytes, self.sparse_cap, _stream())
        return out, self._score_common(launch, nq_local, max_rank, q, span="jaccard_score", fabric=fabric)

    # ------------------------------------------------------------------ K4–K10 from a distance panel
    def rerank_from_panel(self, D: torch.Tensor, key: torch.Tensor, q: int, N: int, k1: int, k2: int, lam: float,
                          detail: dict | None = None, chunk_min: torch.Tensor | None = None,
                          score: tuple | None = None, want_final: bool = True, upper_only: bool = False):
        """D: this rank's panel [n_local, pitch(N)] (row l = probe rank + l*world), key: its row
        maxima.  Returns final[nq_local, pitch(g)] for this rank's queries (use [:, :g]).
        With score = ((q_pids_local, g_pids, q_cams_local, g_cams), max_rank) the last stage is the
        fused Jaccard + mix + CMC/mAP kernel and the return value is (final or None, score results)."""
        cm = self.comm
        r, P = cm.rank, cm.world
        nloc = D.shape[0]
        K = max(k1 + 1, k2)
        if N < K:
            raise NotImplementedError(f"re-ranking needs at least max(k1+1, k2) = {K} samples, got {N}")
        with self.span("select"):
            nbr_local = self.topk_rows(D, N, key, K, chunk_min=chunk_min, upper_only=upper_only)
        with self.span("x_nbr"):
            nbr = cm.gather_rows(nbr_local, N)
        with self.span("krecip"):
            sets = self.krecip_sets(nbr, k1)
            Vc_l, Vv_l, Vn_l = self.krecip_V(sets, k1, D, key, r, P, upper_only=upper_only)
        if P == 1:
            Vc, Vv, Vn = Vc_l, Vv_l, Vn_l
            if k2 != 1:
                with self.span("qe"):
                    Qc, Qv, qoff, Qn, nnz_bound = self.query_expansion(nbr, k2, Vc, Vv, Vn, nloc, r, P)
        else:
            # The ELL rows are (k1+1)(h+1) wide but hold ~k1 entries: only a prefix of `cw` entries per row is exchanged,
            # count, columns and values packed into one byte record per row (one all-gather for the whole of V).  cw is
            # remembered from the previous call; the widest row of THIS call comes back with the pool sizes below (the
            # step's single host read), and if it did not fit the exchange is repeated with the right width.
            cw = min(self._cw_hint.get(k1, 64), Vc_l.shape[1])
            replicate = k2 != 1 and N * k2 <= self.qe_replicate_below
            while True:
                with self.span("x_V"):
                    rec = torch.cat([Vn_l.view(torch.uint8).reshape(nloc, 4),
                                     Vc_l[:, :cw].contiguous().view(torch.uint8).reshape(nloc, 4 * cw),
                                     Vv_l[:, :cw].contiguous().view(torch.uint8).reshape(nloc, 2 * cw)], dim=1)
                    allrec = cm.interleave(cm.gather_stack(cm.pad_rows(rec, N)), N)          # [N, 4 + 6 cw] view
                    Vn = allrec[:, :4].contiguous().view(torch.int32).reshape(N)
                    Vc = allrec[:, 4: 4 + 4 * cw].contiguous().view(torch.int32)
                    Vv = allrec[:, 4 + 4 * cw:].contiguous().view(torch.float16)
                    Vn_fit = torch.clamp(Vn, max=cw)          # keeps a too-narrow attempt inside its buffers
                if replicate:
                    # small problems: every rank expands all N rows itself (0.1 ms at Duke scale) instead of two more
                    # latency-bound all-gathers
                    with self.span("qe"):
                        Qc, Qv, qoff, Qn, nnz_bound = self.query_expansion(nbr, k2, Vc, Vv, Vn_fit, N, 0, 1,
                                                                           extra=Vn.max())
                    widest = self.last_extra
                elif k2 != 1:
                    Qc, Qv, qoff, Qn, nnz_bound, widest = self._query_expansion_sharded(nbr, k2, Vc, Vv, Vn_fit, Vn, N)
                else:
                    widest = int(Vn.max().item())
                self._cw_hint[k1] = min(Vc_l.shape[1], (widest + 7) // 8 * 8 + 8)
                if widest <= cw:
                    break
                cw = self._cw_hint[k1]
            Vn = Vn_fit
        if k2 == 1:
            cap1 = Vc.shape[1]
            qoff = self._empty((N + 1,), torch.int64)
            _lib.call("reid_ell_offsets", N, cap1, _ptr(qoff), _stream())
            Qc, Qv, Qn = Vc.reshape(-1), Vv.reshape(-1), Vn
            nnz_bound = int(Vn.sum().item())
        with self.span("csc"):
            csc = self.inverted_index(Qc, Qv, qoff, Qn, N, nnz_bound)
        nq_local = cm.local_count(q)
        if score is not None:
            assert detail is None
            ids, max_rank = score
            return self.jaccard_score(Qc, Qv, qoff, Qn, csc, D, key, q, N, lam, nq_local, r, P, ids, max_rank,
                                      want_final)
        with self.span("jaccard_mix"):
            res = self.jaccard_mix(Qc, Qv, qoff, Qn, csc, D, key, q, N, lam, nq_local, r, P,
                                   want_jaccard=detail is not None)
        if detail is not None:
            final, J = res
            detail.update(nbr=nbr, sets=sets, Vc=Vc, Vv=Vv, Vn=Vn, Qc=Qc, Qv=Qv, qoff=qoff, Qn=Qn, csc=csc, J=J)
            return final
        return res

    # ------------------------------------------------------------------ sharded symmetric path over peer memory
    def _fabric_for(self, N: int, q: int, K: int, k1: int, k2: int, max_rank: int):
        """(Fabric, Layout) for this problem, or None when peer memory is unavailable (then the NCCL path runs).
        Collective on first use / growth."""
        cm = self.comm
        if cm.world == 1 or not self.use_fabric or self._fabric_failed or self.engine == ENGINES["simt"]:
            return None
        cap1 = v_row_capacity(k1)
        replicate = k2 == 1 or N * k2 <= self.qe_replicate_below
        pool = 0 if replicate else N * k2 * cap1
        if pool * 6 > self.fabric_pool_bytes:
            return None                         # structural pool bound too large for the arena: NCCL path sizes it exactly
        key = (N, q, K, cap1, cm.world, pool, max_rank)
        if key not in self._layouts:
            self._layouts[key] = _fabric.make_layout(N, q, K, cap1, cm.world, pool, max_rank)
        L = self._layouts[key]
        if self.fabric is None:
            self.fabric = _fabric.Fabric(cm, self.device, _lib)
        if not self.fabric.ensure(L.total):
            self._fabric_failed = True
            if cm.rank == 0:
                import warnings
                warnings.warn("reid_b200: peer-memory fabric unavailable (" + self.fabric.last_error
                              + "); using the NCCL all-gather path")
            return None
        return self.fabric, L

    def _tiles_for(self, N: int):
        key = (N, self.comm.rank, self.comm.world)
        if key not in self._tile_cache:
            t = _fabric.tile_list(N, self.comm.rank, self.comm.world)
            self._tile_cache[key] = (torch.from_numpy(t).to(self.device), int(t.shape[0]))
        return self._tile_cache[key]

    def _own_index(self, n: int, rank: int | None = None) -> torch.Tensor:
        """Global index of every local row of `rank` (default: this rank) in an n-row table, on the device."""
        r = self.comm.rank if rank is None else rank
        key = (n, r, self.comm.world)
        if key not in self._own_cache:
            self._own_cache[key] = torch.from_numpy(_fabric.own_row_index(n, r, self.comm.world)).to(self.device)
        return self._own_cache[key]

    def rerank_features_fabric(self, fab, L, feats: torch.Tensor, q: int, k1: int, k2: int, lam: float, normalize: bool,
                               score: tuple | None = None, want_final: bool = True, feats_aug=None):
        """reranking.py:29-100 (+ evaluate.py:40-88 when `score`) on world > 1 ranks with the SYMMETRIC contraction
        sharded by tile rows (snake ownership, fabric.py) and every exchange written into the peers' arenas by kernels:

          prep (replicated) -> upper-triangle tiles of the own tile rows -> per-row bounds -> [stats] -> thresholds,
          scatter selection over the stored tiles -> local top-K records to the row owners [inbox] -> final merge,
          neighbour lists written to every rank [nbr] -> reciprocal sets (replicated) -> V rows of the own probes
          (mirrored distances read from the owner's panel over NVLink) [V] -> query expansion (own rows [pool], or
          replicated for small N) -> inverted index (replicated) -> Jaccard + lambda-mix (+ CMC/mAP) of the own queries
          [scores].  [x] = one peer-memory exchange + one flag barrier; no host synchronisation, no NCCL call.
        Returns what rerank_features returns; rows are this rank's queries in snake order (Engine._own_index(q))."""
        cm = self.comm
        r, P = cm.rank, cm.world
        N, K = feats.shape[0], L.K
        cap1 = L.cap1
        nloc = _fabric.owned_rows(N, r, P)
        st = _stream
        peers = C.c_void_p(fab.peers.data_ptr())
        with self.span("prep"):
            allf = self.prepare(feats, normalize, feats_aug)
        tiles, ntiles = self._tiles_for(N)
        self.last_tiles = ntiles
        panel = fab.view(L.panel, (L.nloc_max, L.ld), torch.float32)
        stats_mine = fab.view(L.stats + r * L.stats_slot, (2, N), torch.int32)
        cpitch = self.lib.reid_chunk_pitch(N)
        ntl = _fabric.owned_tiles(N, r, P)
        pitch_m = max(8, 8 * ntl)
        with self.span("dist"):
            cm_d = self._empty((max(nloc, 1), cpitch), torch.float32)
            cm_m = self._empty((N, pitch_m), torch.float32)
            _lib.call("reid_rowmax_init", _ptr(stats_mine[0]), N, st())
            _lib.call("reid_dist_sharded", _ptr(allf.hi), _ptr(allf.lo), allf.hi.stride(0), _ptr(allf.sq), _ptr(allf.amax),
                      N, allf.dp, _ptr(tiles), ntiles, P, _ptr(panel), L.ld, nloc, _ptr(stats_mine[0]), _ptr(cm_d),
                      cpitch, _ptr(cm_m), pitch_m, st())
        with self.span("select"):
            _lib.call("reid_shard_tau", N, r, P, _ptr(cm_d), cpitch, _ptr(cm_m), pitch_m, K, _ptr(stats_mine[1]), st())
        with self.span("x_stats"):
            fab.bcast(L.stats + r * L.stats_slot, L.stats_slot)
            fab.barrier()
        with self.span("select"):
            cap = 256
            key_g = self._empty((N,), torch.int32)
            key_l = self._empty((max(nloc, 1),), torch.int32)
            dstar = self._empty((N,), torch.float32)
            count = self._empty((N,), torch.int32)
            cthr = self._empty((cpitch,), torch.int32)
            pbuf = self._empty((N, cap), torch.int64)
            stats_all = fab.view(L.stats, (P * L.stats_slot // 4,), torch.int32)
            with self.span(".sel_dstar"):
                _lib.call("reid_shard_dstar", N, r, P, _ptr(stats_all), L.stats_slot // 4, _ptr(key_g), _ptr(key_l),
                          _ptr(dstar), _ptr(count), _ptr(cthr), st())
            with self.span(".sel_push"):
                _lib.call("reid_shard_push", _ptr(panel), nloc, r, P, N, L.ld, _ptr(cm_d), cpitch, _ptr(dstar), _ptr(cthr),
                          _ptr(count), _ptr(pbuf), cap, st())
            with self.span(".sel_merge_local"):
                _lib.call("reid_shard_merge_local", _ptr(panel), L.ld, N, r, P, _ptr(cm_d), cpitch, _ptr(cm_m), pitch_m,
                          _ptr(key_g), _ptr(dstar), _ptr(count), _ptr(pbuf), cap, K, peers, L.inbox, L.nloc_max, st())
        with self.span("x_inbox"):
            fab.barrier()
        with self.span("select"):
            inbox = fab.view(L.inbox, (P, L.nloc_max, K), torch.int64)
            _lib.call("reid_shard_merge_final", nloc, r, P, K, _ptr(inbox), L.nloc_max, peers, L.nbr, None, st())
        with self.span("x_nbr"):
            fab.barrier()
        nbr = fab.view(L.nbr, (N, K), torch.int32)
        Vn = fab.view(L.vn, (N,), torch.int32)
        Vc = fab.view(L.vc, (N, cap1), torch.int32)
        Vv = fab.view(L.vv, (N, cap1), torch.float16)
        with self.span("krecip"):
            with self.span(".krecip_sets"):
                sets = self.krecip_sets(nbr, k1)
            rfull, nfull, rhalf, nhalf = sets
            panels = fab.peers_at(L.panel)
            _lib.call("reid_krecip_V_sharded", _ptr(rfull), _ptr(nfull), _ptr(rhalf), _ptr(nhalf), r, P, nloc, k1,
                      _ptr(panel), L.ld, _ptr(key_l), _ptr(panels), _ptr(Vc), _ptr(Vv), _ptr(Vn), st())
        with self.span("x_V"):
            fab.bcast_ell(L.vn, L.vc, L.vv, cap1, N)
            fab.barrier()
        if k2 == 1:
            qoff = self._empty((N + 1,), torch.int64)
            _lib.call("reid_ell_offsets", N, cap1, _ptr(qoff), st())
            Qc, Qv, Qn = Vc.reshape(-1), Vv.reshape(-1), Vn
            nnz_bound = N * cap1
        elif L.pool_elems == 0:
            with self.span("qe"):                 # small problem: every rank expands all rows itself, no exchange
                Qc, Qv, qoff, Qn, nnz_bound = self.query_expansion(nbr, k2, Vc, Vv, Vn, N, 0, 1)
        else:
            with self.span("qe"):
                ub = self._empty((N,), torch.int32)
                _lib.call("reid_qe_bound", _ptr(nbr), K, k2, _ptr(Vn), 0, 1, N, _ptr(ub), st())
                qoff = self._empty((N + 1,), torch.int64)
                _lib.call("reid_exclusive_scan", _ptr(ub), N, _ptr(qoff), st())
                Qn = fab.view(L.qn, (N,), torch.int32)
                Qc = fab.view(L.qc, (L.pool_elems,), torch.int32)
                Qv = fab.view(L.qv, (L.pool_elems,), torch.float16)
                cap_sort = self._qe_cap_hint()
                if not self._capturing:
                    self._ubmax_pinned.copy_(ub.max().to(torch.int64).reshape(1), non_blocking=True)
                    self._ubmax_event.record()
                    self._ubmax_pending = True
                _lib.call("reid_qe_k2_capped", _ptr(nbr), K, k2, _ptr(Vc), _ptr(Vv), _ptr(Vn), cap1, r, -P, nloc,
                          _ptr(qoff), cap_sort, _ptr(Qc), _ptr(Qv), _ptr(Qn), st())
            with self.span("x_qe"):
                fab.bcast_pool(L.qn, L.qc, L.qv, qoff, N)
                fab.barrier()
            nnz_bound = L.pool_elems
        with self.span("csc"):
            csc = self.inverted_index(Qc, Qv, qoff, Qn, N, nnz_bound)
        nq_local = _fabric.owned_rows(q, r, P)
        if score is not None:
            ids, max_rank = score
            return self.jaccard_score(Qc, Qv, qoff, Qn, csc, panel, key_l, q, N, lam, nq_local, r, -P, ids, max_rank,
                                      want_final, fabric=(fab, L))
        with self.span("jaccard_mix"):
            return self.jaccard_mix(Qc, Qv, qoff, Qn, csc, panel, key_l, q, N, lam, nq_local, r, -P)

    def rerank_features(self, feats: torch.Tensor, q: int, k1: int, k2: int, lam: float, normalize: bool,
                        detail: dict | None = None, score: tuple | None = None, want_final: bool = True,
                        feats_aug: torch.Tensor | None = None):
        """feats: [N, d] float32 on the device, queries first (reranking.py:29-44 then the rest).  feats_aug: optional
        second embedding pass; the rows used are (feats + feats_aug) / 2 (the callers' flip-TTA average, fused)."""
        cm = self.comm
        N = feats.shape[0]
        if cm.world > 1 and detail is None:
            fl = self._fabric_for(N, q, max(k1 + 1, k2), k1, k2, score[1] if score is not None else 50)
            if fl is not None:
                self.last_path = "fabric"
                if N < max(k1 + 1, k2):
                    raise NotImplementedError(f"re-ranking needs at least max(k1+1, k2) samples, got {N}")
                return self.rerank_features_fabric(fl[0], fl[1], feats, q, k1, k2, lam, normalize, score=score,
                                                   want_final=want_final, feats_aug=feats_aug)
        self.last_path = "nccl" if cm.world > 1 else "single"
        with self.span("prep"):
            allf = self.prepare(feats, normalize, feats_aug)
            mine = allf if cm.world == 1 else allf.rows(slice(cm.rank, N, cm.world))
        with self.span("dist"):
            tri = False
            if self.engine == ENGINES["simt"]:
                (D, key), cmin = self.dist(mine, allf, METRIC_SQEUCLIDEAN, want_rowmax=True), None
            else:
                # One rank, symmetric schedule, nobody asked to see D: the mirrored half of the matrix is not written
                # at all (half the 4 N^2-byte output stream).  Selection runs in scatter form (every stored element is
                # offered to both rows it belongs to: reid_topk_rows_scatter) and the V weights read the few mirrored
                # entries they need as D[col, row].  Queries only ever need gallery columns, which lie right of their
                # diagonal block.  A win at every size (Duke: contraction -0.10 ms, selection +0.01 ms).
                tri = (self.upper_only and cm.world == 1 and self.symmetric and detail is None
                       and N >= self.upper_only_min_rows)
                D, key, cmin = self.dist(mine, allf, METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True,
                                         upper_only=tri)
        if detail is not None:
            detail.update(D=D, key=key)
        return self.rerank_from_panel(D, key, q, N, k1, k2, lam, detail, chunk_min=cmin, score=score,
                                      want_final=want_final, upper_only=tri)

    def evaluate_reranked(self, feats: torch.Tensor, q: int, pids: torch.Tensor, cams: torch.Tensor, k1: int, k2: int,
                          lam: float, normalize: bool, max_rank: int = 50, feats_aug: torch.Tensor | None = None,
                          resident: bool = False):
        """The evaluator's hot path end to end (evaluate.py:110-132 with reranking=True): features and ids of
        all N samples (queries first) -> (cmc, mAP, n_valid, ap, first_hit); the q x g matrix is never written.
        resident=True: the caller keeps feats / pids / cams in the same device buffers from call to call (a serving or
        benchmark loop) -- small problems are then captured in a CUDA graph after four identical calls and replayed as
        one launch.  Never inferred: a caller whose allocator merely happens to return the same addresses must not
        trigger captures (measured: 3 ms per evaluation lost to repeated captures in the evaluator's e2e path)."""
        if (self.c_entry and self.comm.world == 1 and self.timers is None and feats_aug is None
                and self.engine != ENGINES["simt"] and self.symmetric and self.upper_only and self.scatter_select
                and self.sparse_cap == 8192 and feats.shape[0] >= max(k1 + 1, k2)):
            return self.evaluate_reranked_single_call(feats, q, pids, cams, k1, k2, lam, normalize, max_rank,
                                                      resident=resident)
        N = feats.shape[0]
        if (resident and self.comm.world > 1 and self.graph_below and N < self.graph_below and self.timers is None
                and feats_aug is None and N * k2 * v_row_capacity(k1) * 14 <= self.qe_nosync_bytes):
            res = self._evaluate_reranked_graph(feats, q, pids, cams, k1, k2, lam, normalize, max_rank)
            if res is not None:
                return res
        return self._evaluate_reranked_staged(feats, q, pids, cams, k1, k2, lam, normalize, max_rank, feats_aug)

    def _evaluate_reranked_staged(self, feats, q, pids, cams, k1, k2, lam, normalize, max_rank, feats_aug=None):
        mine = self.own_queries(feats.shape[0], q, k1, k2, max_rank)
        ids = (pids[:q][mine].contiguous(), pids[q:].contiguous(), cams[:q][mine].contiguous(), cams[q:].contiguous())
        _, res = self.rerank_features(feats, q, k1, k2, lam, normalize, score=(ids, max_rank), want_final=False,
                                      feats_aug=feats_aug)
        return res

    def _evaluate_reranked_graph(self, feats, q, pids, cams, k1, k2, lam, normalize, max_rank):
        """Several ranks, small problem: a step of the peer-memory path is ~60 launches of 5-80 us and the host needs
        ~0.8 ms to issue them -- more than the GPUs need to run them (Duke on 8 B200s).  The step has no host read, no
        library collective and argument-free barriers, so after three eager calls with the same buffers and parameters
        it is captured in a CUDA graph per rank and replayed as one launch.  Every rank decides alike (same call
        sequence), and eager and replayed steps issue the same barriers, so ranks may even mix the two.  Returns None
        when the peer-memory path is not the one in use."""
        N = feats.shape[0]
        if self._fabric_for(N, q, max(k1 + 1, k2), k1, k2, max_rank) is None:
            return None
        key = (feats.data_ptr(), feats.stride(0), N, q, feats.shape[1], pids.data_ptr(), cams.data_ptr(), k1, k2,
               float(lam), bool(normalize), max_rank, self.fabric.base)
        ent = self._graphs.get(key)
        if ent is None:
            if not self._graph_due(key):
                return self._evaluate_reranked_staged(feats, q, pids, cams, k1, k2, lam, normalize, max_rank)
            try:
                torch.cuda.current_stream(self.device).synchronize()
                graph = torch.cuda.CUDAGraph()
                self._capturing = True
                n0 = int(self.lib.reid_launch_count())
                try:
                    with torch.cuda.graph(graph):
                        out = self._evaluate_reranked_staged(feats, q, pids, cams, k1, k2, lam, normalize, max_rank)
                finally:
                    self._capturing = False
                n_launch = int(self.lib.reid_launch_count()) - n0
                # (the graph addresses the inputs by pointer and the key holds those pointers: no reference is kept, so a
                #  caller that frees and re-allocates its inputs every step gets the same addresses -- and this graph -- back)
                ent = self._graphs[key] = (graph, out, n_launch)
            except Exception as e:                              # noqa: BLE001 -- eager from now on (same barriers)
                import warnings
                warnings.warn(f"reid_b200: CUDA-graph capture of the multi-rank step failed ({e}); launching eagerly")
                self.graph_below = 0
                return self._evaluate_reranked_staged(feats, q, pids, cams, k1, k2, lam, normalize, max_rank)
        ent[0].replay()
        self.lib.reid_note_replayed_launches(ent[2])
        self.last_path = "fabric-graph"
        return tuple(t.clone() for t in ent[1])

    def _graph_due(self, key) -> bool:
        """Capture only for a caller that presents the SAME buffers and parameters four times in a row (a benchmark or
        serving loop with resident inputs); inputs re-allocated every call never qualify, and at most 8 graphs exist."""
        if key == self._graph_last_key:
            self._graph_run += 1
        else:
            self._graph_last_key, self._graph_run = key, 1
        return self._graph_run >= 4 and len(self._graphs) < 8

    def evaluate_reranked_single_call(self, feats: torch.Tensor, q: int, pids: torch.Tensor, cams: torch.Tensor, k1: int,
                                      k2: int, lam: float, normalize: bool, max_rank: int = 50, resident: bool = False):
        """The same evaluation through the single C-ABI entry reid_eval_reranked (csrc/eval_e2e.cu): one call, one
        cached workspace, no host read inside, ~30 launches issued from C++ instead of ~25 ctypes round trips.  Same
        kernels, same bits as the staged path (tests/test_gpu_api.py).  One rank, tensor-core engine."""
        assert feats.is_cuda and feats.dtype == torch.float32 and feats.stride(1) == 1
        N, d = feats.shape
        need = int(self.lib.reid_eval_workspace_bytes(N, q, d, k1, k2, max_rank))
        if need < 0:
            raise _lib.ReidError(self.lib.reid_last_error().decode())
        if self._c_ws is None or self._c_ws.numel() < need:
            self._c_ws = None
            self._c_ws = self._empty((need,), torch.uint8)
        pids = pids.contiguous()
        cams = cams.contiguous()

        qn = max(q, 1)
        o_map, o_ap = 8 * ((4 * max_rank + 7) // 8), 8 * ((4 * max_rank + 7) // 8) + 8
        o_nv = o_ap + 8 * qn
        o_first, o_end = o_nv + 8, o_nv + 8 + 4 * qn

        def views(base):                 # one packed buffer: cmc f32 | mAP f64 | ap f64[q] | n_valid i32 | first_hit i32[q]
            return (base[: 4 * max_rank].view(torch.float32), base[o_map: o_map + 8].view(torch.float64),
                    base[o_nv: o_nv + 4].view(torch.int32), base[o_ap: o_ap + 8 * qn].view(torch.float64),
                    base[o_first: o_first + 4 * qn].view(torch.int32))

        def outputs():
            base = self._empty((o_end,), torch.uint8)
            return base, views(base)

        def call(out):
            cmc, mAP, nvalid, ap, first = out
            _lib.call("reid_eval_reranked", _ptr(feats), feats.stride(0), N, q, d, _ptr(pids), _ptr(cams), k1, k2,
                      float(lam), 1 if normalize else 0, max_rank, _ptr(self._c_ws), self._c_ws.numel(), _ptr(cmc),
                      _ptr(mAP), _ptr(nvalid), _ptr(ap), _ptr(first), _stream())

        self.last_path = "single-call"
        if resident and self.graph_below and N < self.graph_below:
            # Small problems are launch-bound (Duke: ~30 launches of 5-400 us): the call is a fixed launch sequence with
            # no allocation and no host read, so it is captured ONCE per (buffers, shape, parameters) and replayed as one
            # graph launch.  The outputs of a replay live in the graph's static tensors until the next replay.
            key = (feats.data_ptr(), feats.stride(0), N, q, d, pids.data_ptr(), cams.data_ptr(), k1, k2, float(lam),
                   bool(normalize), max_rank, self._c_ws.data_ptr())
            ent = self._graphs.get(key)
            if ent is None and self._graph_due(key):
                try:
                    base, out = outputs()
                    call(out)                                   # eager once: function attributes, tensor maps, errors
                    torch.cuda.current_stream(self.device).synchronize()
                    graph = torch.cuda.CUDAGraph()
                    n0 = int(self.lib.reid_launch_count())
                    with torch.cuda.graph(graph):
                        call(out)
                    n_launch = int(self.lib.reid_launch_count()) - n0
                    ent = self._graphs[key] = (graph, base, n_launch)     # inputs are addressed by the pointers in the key
                except Exception as e:                          # noqa: BLE001 -- capture unsupported here: eager from now on
                    import warnings
                    warnings.warn(f"reid_b200: CUDA-graph capture of the evaluation failed ({e}); launching eagerly")
                    self.graph_below = 0
                    ent = None
            if ent is not None:
                ent[0].replay()
                self.lib.reid_note_replayed_launches(ent[2])
                self.last_path = "single-call-graph"
                cmc, mAP, nvalid, ap, first = views(ent[1].clone())     # one copy: results survive the next replay
                return cmc, mAP, nvalid, ap[:q], first[:q]
        _, out = outputs()
        call(out)
        cmc, mAP, nvalid, ap, first = out
        return cmc, mAP, nvalid, ap[:q], first[:q]

    def own_queries(self, N: int, q: int, k1: int, k2: int, max_rank: int = 50) -> torch.Tensor:
        """Global indices of the queries whose rows this rank produces in rerank_features / evaluate_reranked for this
        problem, in the order of the returned rows: snake blocks on the peer-memory path, rank + l * world on the NCCL
        path (collective on first use: the fabric is set up here)."""
        cm = self.comm
        if cm.world > 1 and self._fabric_for(N, q, max(k1 + 1, k2), k1, k2, max_rank) is not None:
            return self._own_index(q)
        return torch.arange(cm.rank, q, cm.world, device=self.device)

    # ------------------------------------------------------------------ K11 + K12
    def rank_rows(self, dist: torch.Tensor, g: int) -> torch.Tensor:
        rows = dist.shape[0]
        keyA = self._empty((rows, g), torch.int32)
        keyB = self._empty((rows, g), torch.int32)
        idxA = self._empty((rows, g), torch.int32)
        ranked = self._empty((rows, g), torch.int32)
        _lib.call("reid_rank_rows", _ptr(dist), rows, g, dist.stride(0), _ptr(keyA), _ptr(keyB), _ptr(idxA),
                  _ptr(ranked), _stream())
        return ranked

    def _score_common(self, launch, nq, max_rank, q_total, span="score", fabric=None):
        cm = self.comm
        if fabric is not None and cm.world > 1:
            return self._score_fabric(launch, nq, max_rank, q_total, span, *fabric)
        ap = self._empty((nq,), torch.float64)
        first = self._empty((nq,), torch.int32)
        hist = self._zeros((max_rank,), torch.int32)
        nvalid = self._zeros((1,), torch.int32)
        with self.span(span):
            launch(ap, first, hist, nvalid)
        q_all = nq if q_total is None else q_total
        with self.span("x_score"):
            if cm.world == 1:
                ap_all, first_all = ap, first
            else:
                # one all-gather: AP float64 | first hit int32 | this rank's CMC histogram and n_valid, as bytes
                nm = cm.nmax(q_all)
                rec = torch.cat([cm.pad_rows(ap, q_all).view(torch.uint8), cm.pad_rows(first, q_all).view(torch.uint8),
                                 hist.view(torch.uint8), nvalid.view(torch.uint8)])
                allrec = cm.gather_stack(rec)
                ap_all = cm.interleave(allrec[:, : 8 * nm].contiguous().view(torch.float64), q_all).contiguous()
                first_all = cm.interleave(allrec[:, 8 * nm: 12 * nm].contiguous().view(torch.int32), q_all).contiguous()
                both = allrec[:, 12 * nm:].contiguous().view(torch.int32).sum(0, dtype=torch.int32)
                hist, nvalid = both[:max_rank].contiguous(), both[max_rank:].contiguous()
        cmc = self._empty((max_rank,), torch.float32)
        mAP = self._empty((1,), torch.float64)
        _lib.call("reid_score_finalize", _ptr(ap_all), q_all, _ptr(hist), max_rank, _ptr(nvalid), _ptr(cmc),
                  _ptr(mAP), _stream())
        return cmc, mAP, nvalid, ap_all, first_all

    def _score_fabric(self, launch, nq, max_rank, q_total, span, fab, L):
        """Scores of the own queries (snake order) written straight into this rank's slot of the arena, one copy to the
        peers, one barrier; every rank then reads all slots in global query order."""
        cm = self.comm
        r, P = cm.rank, cm.world
        assert max_rank <= L.max_rank and nq <= L.nq_max and q_total == L.q
        base = L.scores + r * L.score_slot
        ap = fab.view(base, (L.nq_max,), torch.float64)
        first = fab.view(base + 8 * L.nq_max, (L.nq_max,), torch.int32)
        tail = fab.view(base + 12 * L.nq_max, (L.max_rank + 1,), torch.int32)
        tail.zero_()
        hist, nvalid = tail[:max_rank], tail[L.max_rank:]
        with self.span(span):
            launch(ap, first, hist, nvalid)
        with self.span("x_score"):
            fab.bcast(base, L.score_slot)
            fab.barrier()
            key = ("score_gather", L.q, P, L.score_slot, L.nq_max)
            if key not in self._own_cache:
                pos64 = torch.empty((L.q,), dtype=torch.int64, device=self.device)
                for p in range(P):
                    idx = self._own_index(L.q, p)
                    pos64[idx] = p * (L.score_slot // 8) + torch.arange(idx.numel(), device=self.device)
                pos32 = (pos64 // (L.score_slot // 8)) * (L.score_slot // 4) + 2 * L.nq_max + pos64 % (L.score_slot // 8)
                self._own_cache[key] = (pos64, pos32)
            pos64, pos32 = self._own_cache[key]
            ap_all = fab.view(L.scores, (P * L.score_slot // 8,), torch.float64)[pos64]
            first_all = fab.view(L.scores, (P * L.score_slot // 4,), torch.int32)[pos32]
            tails = fab.view(L.scores, (P, L.score_slot // 4), torch.int32)[:, 3 * L.nq_max: 3 * L.nq_max + L.max_rank + 1]
            both = tails.sum(0, dtype=torch.int32)
            hist_all, nvalid_all = both[:max_rank].contiguous(), both[L.max_rank:].contiguous()
        cmc = self._empty((max_rank,), torch.float32)
        mAP = self._empty((1,), torch.float64)
        _lib.call("reid_score_finalize", _ptr(ap_all), L.q, _ptr(hist_all), max_rank, _ptr(nvalid_all), _ptr(cmc),
                  _ptr(mAP), _stream())
        return cmc, mAP, nvalid_all, ap_all, first_all

    def score(self, ranked: torch.Tensor, q_pids, g_pids, q_cams, g_cams, max_rank: int, q_total: int | None = None):
        """K12 on a materialised ranking.  ranked: this rank's [nq_local, g]; q_pids/q_cams: the
        matching local slices (int64).  Returns (cmc float32[max_rank], mAP float64[1], n_valid[1],
        ap[q], first_hit[q]) as device tensors."""
        nq, g = ranked.shape

        def launch(ap, first, hist, nvalid):
            _lib.call("reid_score_rows", _ptr(ranked), nq, g, _ptr(q_pids), _ptr(g_pids), _ptr(q_cams), _ptr(g_cams),
                      max_rank, _ptr(ap), _ptr(first), _ptr(hist), _ptr(nvalid), _stream())
        return self._score_common(launch, nq, max_rank, q_total)

    def evaluate_dist(self, dist: torch.Tensor, g: int, q_pids, g_pids, q_cams, g_cams, max_rank: int,
                      q_total: int | None = None, materialize_ranking: bool = False):
        """K11 + K12 from distance rows [nq_local, >= g].  Default: the fused counting kernel
        (reid_score_dist); materialize_ranking=True goes through the explicit stable ranking
        (reid_rank_rows + reid_score_rows) — same results."""
        if materialize_ranking:
            with self.span("rank"):
                ranked = self.rank_rows(dist, g)
            return self.score(ranked, q_pids, g_pids, q_cams, g_cams, max_rank, q_total)
        nq = dist.shape[0]

        def launch(ap, first, hist, nvalid):
            _lib.call("reid_score_dist", _ptr(dist), nq, g, dist.stride(0), _ptr(q_pids), _ptr(g_pids), _ptr(q_cams),
                      _ptr(g_cams), max_rank, _ptr(ap), _ptr(first), _ptr(hist), _ptr(nvalid), _stream())
        return self._score_common(launch, nq, max_rank, q_total)


_default = None


def default_engine() -> Engine:
    global _default
    if _default is None:
        _default = Engine()
    return _default
