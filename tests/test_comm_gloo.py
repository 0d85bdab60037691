"""Host-side logic of the multi-rank path (engine.Comm) on CPU: two processes over gloo at 127.0.0.1.

The kernels need a B200, the row-ownership arithmetic does not: cyclic ownership, padding, the rank-major ->
global-order interleave of every all-gather, the packed byte records, the locally derived pool layout of the
query-expansion exchange and the packed score reduction are exercised here against plain numpy.  The full
pipeline's 2-rank == 1-rank identity is tests/multi_rank_check.py (GPU)."""
import importlib
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_rows, ret):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        engine = importlib.import_module("multimodal-reid_b200.engine")
        cm = engine.Comm()
        assert (cm.rank, cm.world) == (rank, world)
        for n in n_rows:
            mine = np.arange(rank, n, world)
            assert cm.local_count(n) == mine.size and cm.nmax(n) == -(-n // world)
            # rows tagged with their global index come back in global order
            table = torch.from_numpy(np.stack([mine * 10 + k for k in range(3)], 1).astype(np.int32)).reshape(-1, 3)
            full = cm.gather_rows(table, n)
            assert full.shape == (n, 3) and np.array_equal(full[:, 0].numpy(), np.arange(n) * 10)
            # packed byte records: int32 | float16 per row through one all-gather, split again after the interleave
            cw = 4
            ci = torch.from_numpy((mine[:, None] * 100 + np.arange(cw)).astype(np.int32)).reshape(-1, cw)
            cv = torch.from_numpy((mine[:, None] + np.arange(cw) / 8).astype(np.float16)).reshape(-1, cw)
            cn = torch.from_numpy((mine % 5).astype(np.int32))
            rec = torch.cat([cn.view(torch.uint8).reshape(mine.size, 4),
                             ci.contiguous().view(torch.uint8).reshape(mine.size, 4 * cw),
                             cv.contiguous().view(torch.uint8).reshape(mine.size, 2 * cw)], dim=1)
            allrec = cm.interleave(cm.gather_stack(cm.pad_rows(rec, n)), n)
            gn = allrec[:, :4].contiguous().view(torch.int32).reshape(n)
            gi = allrec[:, 4: 4 + 4 * cw].contiguous().view(torch.int32)
            gv = allrec[:, 4 + 4 * cw:].contiguous().view(torch.float16)
            assert np.array_equal(gn.numpy(), np.arange(n) % 5)
            assert np.array_equal(gi.numpy(), np.arange(n)[:, None] * 100 + np.arange(cw))
            assert np.array_equal(gv.numpy(), (np.arange(n)[:, None] + np.arange(cw) / 8).astype(np.float16))
            # sharded upload: every rank copies its contiguous block of the (ragged) host batches, all-gather restores all rows
            d = 5
            host = torch.arange(n * d, dtype=torch.float32).reshape(n, d)
            cuts = sorted({0, n} | {min(n, c) for c in (3, 4, 11)})
            parts = [host[a:b] for a, b in zip(cuts[:-1], cuts[1:])]
            lo, hi = cm.block_bounds(n)
            assert (lo, hi) == (min(n, rank * cm.nmax(n)), min(n, (rank + 1) * cm.nmax(n)))
            assert torch.equal(cm.upload_sharded(parts, torch.device("cpu")), host)
            # pool layout: every rank derives every rank's pool size and row offsets from the global bounds
            rng = np.random.default_rng(n)
            ub = np.zeros(cm.nmax(n) * world, np.int32)
            ub[:n] = rng.integers(0, 50, n)
            totals, excl = cm.pool_layout(torch.from_numpy(ub), world)
            for p in range(world):
                own = ub[p::world]
                assert int(totals[p]) == own.sum()
                assert np.array_equal(excl[p].numpy(), np.cumsum(own) - own)
            # global offsets as the engine builds them address rank-major pools of pitch `pitch`
            pitch = int(totals.max()) + cm.nmax(n)
            base = torch.arange(world, dtype=torch.int64).unsqueeze(1) * pitch
            qoff = cm.interleave(excl + base, n).numpy()
            for i in range(n):
                p, l = i % world, i // world
                assert qoff[i] == p * pitch + (np.cumsum(ub[p::world]) - ub[p::world])[l]
            # scores: AP (float64) | first hit (int32) | this rank's histogram and n_valid, one byte record per rank
            ap = torch.from_numpy(mine.astype(np.float64) / 7)
            first = torch.from_numpy(mine.astype(np.int32) * 3)
            hist = torch.full((5,), rank + 1, dtype=torch.int32)
            nvalid = torch.tensor([mine.size], dtype=torch.int32)
            r = torch.cat([cm.pad_rows(ap, n).view(torch.uint8), cm.pad_rows(first, n).view(torch.uint8),
                           hist.view(torch.uint8), nvalid.view(torch.uint8)])
            allr = cm.gather_stack(r)
            nm = cm.nmax(n)
            ap_all = cm.interleave(allr[:, : 8 * nm].contiguous().view(torch.float64), n)
            first_all = cm.interleave(allr[:, 8 * nm: 12 * nm].contiguous().view(torch.int32), n)
            assert np.array_equal(ap_all.numpy(), np.arange(n) / 7) and np.array_equal(first_all.numpy(), np.arange(n) * 3)
            both = allr[:, 12 * nm:].contiguous().view(torch.int32).sum(0, dtype=torch.int32)
            assert both[:5].tolist() == [sum(range(1, world + 1))] * 5 and int(both[5]) == n
            chk = torch.tensor([rank + 1], dtype=torch.int32)
            cm.all_reduce_sum(chk)
            assert int(chk) == sum(range(1, world + 1))
            assert cm.max_int(rank * 5 + n, "cpu") == (world - 1) * 5 + n
        dist.barrier()
        dist.destroy_process_group()
        ret[rank] = "ok"
    except Exception as e:          # noqa: BLE001 - reported to the parent
        import traceback
        ret[rank] = traceback.format_exc() + repr(e)


def test_comm_two_ranks_gloo():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    procs = [mp.get_context("spawn").Process(target=_worker, args=(r, world, port, (2, 7, 64, 1001), ret))
             for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert dict(ret) == {0: "ok", 1: "ok"}, dict(ret)


def test_comm_single_rank_is_identity():
    sys.path.insert(0, ROOT)
    engine = importlib.import_module("multimodal-reid_b200.engine")
    cm = engine.Comm(single=True)
    t = torch.arange(12).reshape(4, 3)
    assert cm.gather_rows(t, 4) is t and cm.local_count(9) == 9 and cm.nmax(9) == 9
    assert torch.equal(cm.interleave(cm.gather_stack(t), 4), t)
    totals, excl = cm.pool_layout(torch.tensor([3, 0, 5], dtype=torch.int32), 1)
    assert totals.tolist() == [8] and excl.tolist() == [[0, 3, 3]]
