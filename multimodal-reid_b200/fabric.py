"""Peer-memory fabric of a multi-rank evaluation (one process per GPU on one NVSwitch box).

Host side of csrc/fabric.cu: the snake-block ownership arithmetic, the tile list of a rank's share of the symmetric
contraction, the arena layout every rank derives identically, and the `Fabric` object (arena allocation, CUDA-IPC
exchange of the handles, the device array of peer base pointers, the stream-ordered flag barrier).  torch.distributed
is used ONCE, to exchange the 64-byte IPC handles; after that no library collective runs inside an evaluation.
"""
from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

TILE = 256


# ----------------------------------------------------------------------------------------------------------------------
# ownership arithmetic (mirrors csrc/common.cuh; pure Python so the CPU tests can check it)
# ----------------------------------------------------------------------------------------------------------------------
def snake_owner(t: int, P: int) -> int:
    pos = t % (2 * P)
    return pos if pos < P else 2 * P - 1 - pos


def snake_local_tile(t: int, P: int) -> int:
    pos = t % (2 * P)
    return 2 * (t // (2 * P)) + (1 if pos >= P else 0)


def snake_global_tile(lt: int, r: int, P: int) -> int:
    return 2 * P * (lt >> 1) + ((2 * P - 1 - r) if (lt & 1) else r)


def snake_tiles_below(t_end: int, r: int, P: int) -> int:
    pos = t_end % (2 * P)
    return 2 * (t_end // (2 * P)) + (1 if pos > r else 0) + (1 if pos > 2 * P - 1 - r else 0)


def n_tiles(n: int) -> int:
    return (n + TILE - 1) // TILE


def owned_tiles(n: int, r: int, P: int) -> int:
    return snake_tiles_below(n_tiles(n), r, P)


def owned_rows(n: int, r: int, P: int) -> int:
    """Rows of an n-row table that rank r owns (its local row count; only the global last tile row can be partial)."""
    T = n_tiles(n)
    nloc = snake_tiles_below(T, r, P) * TILE
    if T > 0 and snake_owner(T - 1, P) == r:
        nloc -= T * TILE - n
    return nloc


def owned_rows_max(n: int, P: int) -> int:
    return max(owned_rows(n, r, P) for r in range(P))


def own_row_index(n: int, r: int, P: int) -> np.ndarray:
    """Global row of every local row of rank r, in local order (int64)."""
    out = np.empty(owned_rows(n, r, P), dtype=np.int64)
    at = 0
    for lt in range(owned_tiles(n, r, P)):
        r0 = snake_global_tile(lt, r, P) * TILE
        r1 = min(n, r0 + TILE)
        out[at: at + r1 - r0] = np.arange(r0, r1)
        at += r1 - r0
    assert at == out.size
    return out


RASTER_GROUP = 8


def tile_list(n: int, r: int, P: int) -> np.ndarray:
    """(tile row, tile column) of every upper-triangle tile rank r computes, int32 [ntiles, 2].  Order: groups of
    RASTER_GROUP owned tile rows, column by column inside a group -- the tiles in flight at any moment share a few A
    row blocks and neighbouring B column blocks (L2 reuse), like the single-rank raster of dist_tc2.cu."""
    T = n_tiles(n)
    mine = [snake_global_tile(lt, r, P) for lt in range(owned_tiles(n, r, P))]
    out = []
    for g0 in range(0, len(mine), RASTER_GROUP):
        grp = mine[g0: g0 + RASTER_GROUP]
        for nb in range(grp[0], T):
            for mb in grp:
                if mb <= nb:
                    out.append((mb, nb))
    return np.asarray(out, dtype=np.int32).reshape(-1, 2)


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


@dataclasses.dataclass(frozen=True)
class Layout:
    """Byte offsets inside the arena; every rank computes the same object from the same arguments."""
    n: int
    q: int
    K: int
    cap1: int
    world: int
    nloc_max: int
    nq_max: int
    ld: int
    pool_elems: int
    max_rank: int
    flags: int
    stats: int
    stats_slot: int
    inbox: int
    nbr: int
    vn: int
    vc: int
    vv: int
    qn: int
    qc: int
    qv: int
    scores: int
    score_slot: int
    panel: int
    total: int


def make_layout(n: int, q: int, K: int, cap1: int, world: int, pool_elems: int, max_rank: int = 50) -> Layout:
    nloc_max = _align(owned_rows_max(n, world), TILE)           # whole tile rows: the panel is written by 32 x 32 boxes
    nq_max = max(1, owned_rows_max(q, world))
    ld = (n + 31) // 32 * 32
    off = 0

    def take(nbytes):
        nonlocal off
        at = off
        off = _align(off + nbytes)
        return at

    flags = take(4 * 64)
    stats_slot = _align(2 * n * 4, 16)
    stats = take(world * stats_slot)
    inbox = take(world * nloc_max * K * 8)
    nbr = take(n * K * 4)
    vn = take(n * 4)
    vc = take(n * cap1 * 4)
    vv = take(n * cap1 * 2)
    qn = take(n * 4)
    qc = take(pool_elems * 4)
    qv = take(pool_elems * 2)
    score_slot = _align(nq_max * 12 + (max_rank + 1) * 4, 16)   # AP float64 | first hit int32 | CMC histogram | n_valid
    scores = take(world * score_slot)
    panel = take(nloc_max * ld * 4)
    return Layout(n, q, K, cap1, world, nloc_max, nq_max, ld, pool_elems, max_rank, flags, stats, stats_slot, inbox, nbr,
                  vn, vc, vv, qn, qc, qv, scores, score_slot, panel, off)


# ----------------------------------------------------------------------------------------------------------------------
class _Cai:
    """Minimal __cuda_array_interface__ holder so that torch can alias arena memory without owning it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
                                         "strides": None}


class Fabric:
    """One arena per rank, mapped by every peer.  Collective calls (ensure) must be made by all ranks together."""

    def __init__(self, comm, device, lib_mod):
        import torch
        self.torch = torch
        self.comm = comm
        self.rank, self.world = comm.rank, comm.world
        self.device = device
        self._lib = lib_mod
        self.lib = lib_mod.load()
        self.nbytes = 0
        self.base = 0                   # own arena (device address)
        self.peer_ptrs = []             # arena base of every rank as mapped in THIS process
        self.peers = None               # int64 device tensor [world]: the same, for the kernels
        self.arena = None               # torch uint8 view of the own arena
        self.epoch = 0
        self.last_error = ""

    # -- lifecycle --------------------------------------------------------------------------------------------------
    def ensure(self, nbytes: int) -> bool:
        """Collective: (re)creates the arenas when the current ones are smaller than nbytes.  Returns False on EVERY rank
        when any rank could not allocate, export or map (no CUDA IPC between these processes, no peer access): the
        caller then uses the NCCL path.  Local failures never skip a collective, so no rank is left waiting."""
        torch = self.torch
        if nbytes <= self.nbytes:
            return True
        self.release()
        nbytes = _align(int(nbytes * 1.05), 1 << 21)
        ok, why = True, ""
        handle = (C.c_ubyte * 64)()
        try:
            ptr = C.c_void_p()
            self._lib.call("reid_fabric_alloc", nbytes, C.byref(ptr))
            self.base, self.nbytes = int(ptr.value), nbytes
            self.arena = torch.as_tensor(_Cai(self.base, nbytes), device=self.device)
            self.arena[:4096].zero_()                                      # the barrier flags (first bytes) start at 0
            self._lib.call("reid_fabric_export", C.c_void_p(self.base), handle)
        except Exception as e:                                             # noqa: BLE001 -- reported below, collectively
            ok, why = False, f"{type(e).__name__}: {e}"
        # the one library collective of the fabric: 64-byte handles, as Python objects (works over NCCL and gloo alike)
        allh = [None] * self.world
        self.comm.dist.all_gather_object(allh, bytes(handle) if ok else b"", group=self.comm.group)
        ok = ok and all(len(h) == 64 for h in allh)
        self.peer_ptrs = []
        if ok:
            try:
                for p in range(self.world):
                    if p == self.rank:
                        self.peer_ptrs.append(self.base)
                        continue
                    h = (C.c_ubyte * 64)(*allh[p])
                    out = C.c_void_p()
                    self._lib.call("reid_fabric_open", h, C.byref(out))
                    self.peer_ptrs.append(int(out.value))
                self.peers = torch.tensor(self.peer_ptrs, dtype=torch.int64, device=self.device)
            except Exception as e:                                         # noqa: BLE001
                ok, why = False, f"{type(e).__name__}: {e}"
        if ok:
            torch.cuda.synchronize(self.device)                            # flags zeroed before anybody signals
        oks = [None] * self.world
        self.comm.dist.all_gather_object(oks, (bool(ok), why), group=self.comm.group)   # also: every arena is mapped
        self.epoch = 0
        if not all(o[0] for o in oks):
            self.last_error = "; ".join(f"rank {p}: {o[1]}" for p, o in enumerate(oks) if not o[0])
            self._teardown_local()
            return False
        return True

    def _teardown_local(self):
        for p, ptr in enumerate(self.peer_ptrs):
            if p != self.rank and ptr:
                self.lib.reid_fabric_close(C.c_void_p(ptr))
        self.arena = None
        if self.base:
            self.lib.reid_fabric_free(C.c_void_p(self.base))
        self.base, self.nbytes, self.peer_ptrs, self.peers = 0, 0, [], None

    def release(self):
        if not self.base:
            return
        self.torch.cuda.synchronize(self.device)
        self.comm.dist.barrier(group=self.comm.group)                     # nobody is still reading a peer's arena
        self._teardown_local()
        self.comm.dist.barrier(group=self.comm.group)

    # -- views and primitives -------------------------------------------------------------------------------------
    def view(self, offset: int, shape, dtype):
        torch = self.torch
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        return self.arena[offset: offset + n].view(dtype).view(*shape)

    def peers_at(self, offset: int):
        """Device int64 array: address of `offset` inside every rank's arena."""
        return self.peers + offset

    def _ptr(self):
        return C.c_void_p(self.peers.data_ptr())

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def barrier(self, flags_off: int = 0):
        # epoch 0: the kernel counts its own barriers in device memory (no per-call argument: graph-capturable)
        self._lib.call("reid_fabric_barrier", self._ptr(), flags_off, self.rank, self.world, 0, self._stream())

    def bcast(self, offset: int, nbytes: int):
        self._lib.call("reid_fabric_bcast", self._ptr(), self.rank, self.world, offset, _align(nbytes, 16), self._stream())

    def bcast_rows(self, offset: int, row_bytes: int, n: int):
        self._lib.call("reid_fabric_bcast_rows", self._ptr(), self.rank, self.world, offset, row_bytes, n, self._stream())

    def bcast_ell(self, off_cnt: int, off_cols: int, off_vals: int, cap1: int, n: int):
        self._lib.call("reid_fabric_bcast_ell", self._ptr(), self.rank, self.world, off_cnt, off_cols, off_vals, cap1, n,
                       self._stream())

    def bcast_pool(self, off_cnt: int, off_cols: int, off_vals: int, qoff, n: int):
        self._lib.call("reid_fabric_bcast_pool", self._ptr(), self.rank, self.world, off_cnt, off_cols, off_vals,
                       C.c_void_p(qoff.data_ptr()), n, self._stream())
