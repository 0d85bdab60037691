"""One small pass through EVERY kernel family for compute-sanitizer (memcheck / racecheck / synccheck), SURVEY 5:

    compute-sanitizer --tool memcheck python tests/sanitizer_smoke.py

The tiny shape (60 x 300 x 64) goes through the prologue (incl. the flip-TTA form), the tcgen05 contraction in its full,
mirrored and upper-only schedules and in streamed blocks, both selection forms, reciprocal sets, V weights, query
expansion, the inverted index, the sparse and the dense Jaccard kernels, all scoring kernels, the single-call entry, the
streamed retrieval / CMC-mAP sweeps, the no-dense-matrix re-ranking, and the training-side ops.  Results are compared
with the staged path so that a sanitizer-induced reordering that changed a bit would be seen."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    engine = importlib.import_module("multimodal-reid_b200.engine")
    synth = importlib.import_module("multimodal-reid_b200.synth")
    losses = importlib.import_module("multimodal-reid_b200.losses")
    eng = engine.Engine()
    dev = eng.device
    sh = synth.SHAPES["tiny"]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 0)
    X = torch.cat([qf, gf]).to(dev)
    pids, cams = torch.cat([qp, gp]).to(dev), torch.cat([qc, gc]).to(dev)
    q, g = sh.q, sh.g
    ids = (pids[:q].contiguous(), pids[q:].contiguous(), cams[:q].contiguous(), cams[q:].contiguous())
    # one-rank default path: C entry and staged path
    a = eng.evaluate_reranked(X, q, pids, cams, 20, 6, 0.3, True)
    eng.c_entry = False
    b = eng.evaluate_reranked(X, q, pids, cams, 20, 6, 0.3, True)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    # flip-TTA prologue, full / mirrored schedules, pull-form selection, dense Jaccard, matrix output, explicit ranking
    eng.prepare(X, True, X.flip(0).contiguous())
    for sym, scatter, cap in ((False, True, 8192), (True, False, 0)):
        eng.symmetric, eng.scatter_select, eng.sparse_cap = sym, scatter, cap
        det = {}
        final = eng.rerank_features(X, q, 20, 6, 0.3, True, detail=det)
        r1 = eng.evaluate_dist(final, g, *ids, 50)
        r2 = eng.evaluate_dist(final, g, *ids, 50, materialize_ranking=True)
        assert torch.equal(r1[0], r2[0])
    eng.symmetric, eng.scatter_select, eng.sparse_cap = True, True, 8192
    final2 = eng.rerank_features(X, q, 7, 1, 0.3, True)               # k2 == 1 (ELL offsets), upper-only + sparse Jaccard
    # streamed host path (blocks), cosine metric
    eng.evaluate_reranked_host([X.cpu().pin_memory()], q, pids, cams, 20, 6, 0.3, 50)
    f = eng.prepare(X, False)
    eng.dist(f.rows(slice(0, q)), f.rows(slice(q, q + g)), engine.METRIC_ARCCOS)
    # streamed retrieval, streamed CMC/mAP, re-ranking without a dense matrix
    allf = eng.prepare(X, True)
    eng.evaluate_streamed(allf.rows(slice(0, q)), allf.rows(slice(q, q + g)), *ids, 50, 10, rows_per_pass=256, cols_per_pass=256)
    eng.evaluate_streamed_reranked(allf, q, pids, cams, 20, 6, 0.3, topk=20, rows_per_pass=256, cols_per_pass=256, cap_o=4096)
    # training-side ops, forward + backward
    x = torch.randn(64, 128, device=dev, requires_grad=True)
    lab = torch.arange(64, device=dev) // 4
    loss = losses.WeightedRegularizedTriplet(0.3)(x, lab)
    (loss[0] if isinstance(loss, tuple) else loss).backward()
    torch.cuda.synchronize()
    print("sanitizer smoke ok: mAP", float(a[1]))


if __name__ == "__main__":
    main()
