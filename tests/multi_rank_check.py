"""Run under torchrun (one rank per GPU; or gloo on CPU for the host-side logic only):
   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multi_rank_check.py

Asserts that the row-sharded pipeline returns, on every rank, exactly the bits a single rank
computes alone (SURVEY.md §8e invariance requirement): the re-ranked rows of this rank's queries,
the per-query APs, the CMC curve and mAP."""
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    # REID_MR_SAME_GPU=1: every rank on cuda:0, rendezvous over gloo -- the peer-memory path (CUDA IPC arenas, flag
    # barriers, sharded symmetric contraction) is exercised functionally on a ONE-GPU box; the contexts time-slice, so a
    # barrier costs a time slice, and the NCCL path (which refuses two ranks on one device) is skipped
    same_gpu = os.environ.get("REID_MR_SAME_GPU", "0") == "1"
    if same_gpu:
        local = 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if same_gpu:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    engine = importlib.import_module("multimodal-reid_b200.engine")
    synth = importlib.import_module("multimodal-reid_b200.synth")
    multi = engine.Engine(dev)
    single = engine.Engine(dev, comm=engine.Comm(single=True))
    single.symmetric = False        # NCCL path: every rank runs the full tile schedule on its cyclic rows
    single_sym = engine.Engine(dev, comm=engine.Comm(single=True))   # peer-memory path == the one-rank DEFAULT (symmetric
    #                                                                  upper-only contraction): mirrored entries are copies
    assert multi.comm.world == world and single.comm.world == 1
    cases = [("small", 20, 6, 0), ("small", 50, 15, 1), ("mid", 20, 6, 2), ("tiny", 7, 1, 3)]
    if os.environ.get("REID_MR_BIG", "0") == "1":
        cases.append(("duke", 20, 6, 4))                 # 78 tile rows: every rank owns several, mirrored reads cross ranks
    for name, k1, k2, seed in cases:
        sh = synth.SHAPES[name]
        qf, gf, qp, gp, qc, gc = synth.make(sh, seed)
        feats = torch.cat([qf, gf]).to(dev)
        q, g = sh.q, sh.g
        N = q + g
        ids = [t.to(dev) for t in (qp, gp, qc, gc)]
        pids, cams = torch.cat([qp, gp]).to(dev), torch.cat([qc, gc]).to(dev)
        refs = {}
        for nm, eng1 in (("fabric", single_sym), ("nccl", single)):
            ref_ = eng1.rerank_features(feats, q, k1, k2, 0.3, normalize=True)
            refs[nm] = (ref_, eng1.evaluate_dist(ref_, g, *ids, 50),
                        eng1.evaluate_reranked(feats, q, pids, cams, k1, k2, 0.3, normalize=True))
        # peer-memory path (symmetric contraction sharded by tile rows) and the NCCL all-gather path (cyclic rows)
        for use_fabric in ((True,) if same_gpu else (True, False)):
            multi.use_fabric = use_fabric
            for below in (300_000, 0):          # query expansion replicated on every rank / sharded + exchanged
                multi.qe_replicate_below = below
                mine = multi.own_queries(N, q, k1, k2)
                got = multi.rerank_features(feats, q, k1, k2, 0.3, normalize=True)
                path = multi.last_path
                assert path == ("fabric" if use_fabric and not multi._fabric_failed else "nccl"), path
                ref, r1, e1 = refs[path]
                assert got.shape[0] == mine.numel()
                assert torch.equal(got[:, :g], ref[mine, :g]), \
                    f"{name}: rank {rank} rows differ from the single-rank result ({path}, qe_replicate_below={below})"
                e2 = multi.evaluate_reranked(feats, q, pids, cams, k1, k2, 0.3, normalize=True)
                for a, b, what in zip(e1, e2, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
                    assert torch.equal(a, b), f"{name}: fused {what} differs on rank {rank} ({path}, below={below})"
            if path == "nccl":
                r2 = multi.evaluate_dist(got, g, ids[0][mine].contiguous(), ids[1], ids[2][mine].contiguous(), ids[3], 50,
                                         q_total=q)
                for a, b, what in zip(r1, r2, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
                    assert torch.equal(a, b), f"{name}: {what} differs on rank {rank}"
            if rank == 0:
                print(f"{name} k1={k1} k2={k2} [{path}]: {world}-rank result == 1-rank result, mAP {float(e2[1]):.6f}",
                      flush=True)
    multi.use_fabric = True
    multi.qe_replicate_below = 300_000
    # small problems on several ranks: after three eager calls with the same buffers the step is captured in a CUDA graph
    # per rank and replayed as one launch (argument-free barriers) -- same bits, call after call
    if not multi._fabric_failed:
        for name, k1, k2, seed in (("small", 20, 6, 0), ("mid", 20, 6, 2)):
            sh = synth.SHAPES[name]
            qf, gf, qp, gp, qc, gc = synth.make(sh, seed)
            feats = torch.cat([qf, gf]).to(dev)
            pids, cams = torch.cat([qp, gp]).to(dev), torch.cat([qc, gc]).to(dev)
            want = single_sym.evaluate_reranked(feats, sh.q, pids, cams, k1, k2, 0.3, normalize=True)
            paths = []
            for it in range(7):
                got = multi.evaluate_reranked(feats, sh.q, pids, cams, k1, k2, 0.3, normalize=True, resident=True)
                paths.append(multi.last_path)
                for a, b, what in zip(want, got, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
                    assert torch.equal(a, b), f"{name}: {what} differs on rank {rank} at call {it} ({multi.last_path})"
            assert paths[-1] == "fabric-graph", paths
            if rank == 0:
                print(f"{name}: {world}-rank graph-replayed step == 1-rank result ({paths})", flush=True)
    # reranking.py:43-44 / :33-34 with a caller-supplied N x N matrix on several ranks: replicas, own rows returned
    rr = importlib.import_module("multimodal-reid_b200.reranking")
    sh = synth.SHAPES["tiny"]
    qf, gf, *_ = synth.make(sh, 8)
    fn = torch.nn.functional.normalize(torch.cat([qf, gf]), dim=1, p=2)
    Lm = torch.rand((fn.shape[0], fn.shape[0]), generator=torch.Generator().manual_seed(3)) * 0.05
    for only in (False, True):
        a = rr.re_ranking_device(fn[: sh.q], fn[sh.q:], 20, 6, 0.3, local_distmat=Lm, only_local=only, eng=multi)
        b = rr.re_ranking_device(fn[: sh.q], fn[sh.q:], 20, 6, 0.3, local_distmat=Lm, only_local=only, eng=single_sym)
        assert torch.equal(a, b[rank::world]), f"local_distmat (only_local={only}) differs on rank {rank}"
    if rank == 0:
        print(f"local_distmat on {world} ranks == 1 rank (additive and only_local)", flush=True)
    if same_gpu:
        assert not multi._fabric_failed, multi.fabric.last_error
        dist.barrier()
        dist.destroy_process_group()
        return
    # the drop-in evaluator on host batches: sharded upload + all-gather == replicated upload, bit for bit
    import contextlib
    import io
    ev = importlib.import_module("multimodal-reid_b200.evaluate")
    sh = synth.SHAPES["mid"]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 5)
    feats, pids, cams = torch.cat([qf, gf]), torch.cat([qp, gp]), torch.cat([qc, gc])
    res = []
    for shard in (True, False):
        engine.default_engine().shard_h2d = shard
        e = ev.R1_mAP_eval(sh.q, feat_norm=True, reranking=True, rerank_params=(20, 6, 0.3))
        e.reset()
        for a in range(0, feats.shape[0], 997):          # ragged batches, like a data loader's last batch
            e.update((feats[a:a + 997], pids[a:a + 997], cams[a:a + 997]))
        with contextlib.redirect_stdout(io.StringIO()):
            res.append(e.compute())
    assert (res[0][0] == res[1][0]).all() and res[0][1] == res[1][1], "sharded upload changed the result"
    if rank == 0:
        print(f"evaluator on host batches: sharded upload == replicated upload, mAP {res[0][1]:.6f}", flush=True)
    # streamed retrieval + CMC/mAP: sharded queries + all-gathered lists == one rank, bit for bit
    retrieval = importlib.import_module("multimodal-reid_b200.retrieval")
    engine.default_engine().shard_h2d = True
    g_ = torch.Generator().manual_seed(9)
    qn, gn = 203, 3001
    gp2 = torch.randint(0, 40, (gn,), generator=g_)
    qp2 = torch.randint(0, 40, (qn,), generator=g_)
    qc2, gc2 = torch.randint(0, 5, (qn,), generator=g_), torch.randint(0, 5, (gn,), generator=g_)
    cent = torch.randn(40, 64, generator=g_)
    qf2 = cent[qp2] + 1.5 * torch.randn(qn, 64, generator=g_)
    gf2 = cent[gp2] + 1.5 * torch.randn(gn, 64, generator=g_)
    cmc_m, map_m, idx_m, val_m = retrieval.streamed_eval(qf2, gf2, qp2, gp2, qc2, gc2, topk=30, rows_per_pass=64,
                                                         cols_per_pass=700)
    allf = single.prepare(torch.cat([qf2, gf2]).to(dev), True)
    ids2 = [t.to(dev) for t in (qp2, gp2, qc2, gc2)]
    (cmc_s, map_s, _, _, _), (idx_s, val_s) = single.evaluate_streamed(
        allf.rows(slice(0, qn)), allf.rows(slice(qn, qn + gn)), ids2[0], ids2[1], ids2[2], ids2[3], 50, 30,
        rows_per_pass=64, cols_per_pass=700)
    assert (cmc_m == cmc_s.cpu().numpy()).all() and map_m == float(map_s.item()), "streamed CMC/mAP differs across ranks"
    assert (idx_m == idx_s.cpu().numpy()).all() and (val_m == val_s.cpu().numpy()).all(), "streamed top-k differs"
    if rank == 0:
        print(f"streamed retrieval + CMC/mAP: {world}-rank result == 1-rank result, mAP {map_m:.6f}", flush=True)
    # k-reciprocal re-ranking with no dense matrix: sharded sweeps + all-gathered lists == one rank
    cm_m, map_m, ridx_m, rval_m = retrieval.streamed_rerank_eval(qf2, gf2, qp2, gp2, qc2, gc2, k1=12, k2=4, lambda_value=0.3,
                                                                 topk=25, rows_per_pass=256, cols_per_pass=768,
                                                                 cap_o=8192)
    pids2, cams2 = torch.cat([ids2[0], ids2[1]]), torch.cat([ids2[2], ids2[3]])
    (cmc_s, map_s, _, _, _), (ridx_s, rval_s), _ = single.evaluate_streamed_reranked(
        allf, qn, pids2, cams2, 12, 4, 0.3, topk=25, rows_per_pass=256, cols_per_pass=768, cap_o=8192)
    assert (cm_m == cmc_s.cpu().numpy()).all() and map_m == float(map_s.item()), "streamed re-ranked CMC/mAP differs"
    assert (ridx_m == ridx_s.cpu().numpy()).all() and (rval_m == rval_s.cpu().numpy()).all(), "re-ranked top-k differs"
    if rank == 0:
        print(f"streamed re-ranking (no dense matrix): {world}-rank result == 1-rank result, mAP {map_m:.6f}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
