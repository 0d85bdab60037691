"""Freezes outputs of the UNMODIFIED reference (imported from /root/reference, never copied) into
tests/golden/*.npz so that parity can be checked where the reference tree is not mounted (the GPU
box).  TEST INFRASTRUCTURE ONLY.  Run in the build container:  python oracle/make_golden.py

Each file holds the exact inputs (already L2-normalised features, ids, cameras), the N x N matrix
the reference's own distance block produces (reranking.py:36-41), and the reference outputs of
  re_ranking(q, g, k1, k2, lam)                                   -> final
  re_ranking(q, g, k1, k2, lam, local_distmat=D, only_local=True) -> final_local   (reranking.py:33-34)
  eval_func(final, ...) / eval_func(euclidean_distance(q, g), ...)                  (evaluate.py:29-88)
  euclidean_distance / cosine_similarity                                            (evaluate.py:7-26)
  R1_mAP_eval(num_query, reranking=False/True).compute() on the raw features        (evaluate.py:91-135)
"""
import contextlib
import importlib
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _refload  # noqa: E402

EU_ROWS = 40   # rows of the q x g matrices kept (file size)
CASES = [  # name, shape, k1, k2, lam, seed
    ("tiny_k20", "tiny", 20, 6, 0.3, 0),
    ("tiny_k50", "tiny", 50, 15, 0.3, 1),
    ("tiny_k5", "tiny", 5, 1, 0.3, 2),
    ("tiny_k25", "tiny", 25, 3, 0.5, 3),
    ("small_k20", "small", 20, 6, 0.3, 0),
]


def main():
    synth = importlib.import_module("multimodal-reid_b200.synth")
    ev, rr = _refload.reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, shape, k1, k2, lam, seed in CASES:
        sh = synth.SHAPES[shape]
        qf, gf, qp, gp, qc, gc = synth.make(sh, seed)
        raw = torch.cat([qf, gf])
        f = torch.nn.functional.normalize(raw, dim=1, p=2)
        qn, gn = f[: sh.q], f[sh.q:]
        N = f.shape[0]
        # the reference's own distance block (reranking.py:36-41), same calls
        D = torch.pow(f, 2).sum(dim=1, keepdim=True).expand(N, N) + \
            torch.pow(f, 2).sum(dim=1, keepdim=True).expand(N, N).t()
        D = D.clone()
        D.addmm_(f, f.t(), beta=1, alpha=-2)
        D = D.numpy()
        with contextlib.redirect_stdout(io.StringIO()):
            final = np.ascontiguousarray(rr.re_ranking(qn, gn, k1, k2, lam))
            final_local = np.ascontiguousarray(rr.re_ranking(qn, gn, k1, k2, lam, local_distmat=D.copy(),
                                                             only_local=True))
            eu = ev.euclidean_distance(qn, gn)
            co = ev.cosine_similarity(qf, gf)
            cmc_rr, map_rr = ev.eval_func(final, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
            cmc_eu, map_eu = ev.eval_func(eu, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
            res = {}
            for flag in (False, True):
                e = ev.R1_mAP_eval(sh.q, max_rank=10, feat_norm=True, reranking=flag)
                e.reset()
                e.update((raw, torch.cat([qp, gp]), torch.cat([qc, gc])))
                res[flag] = e.compute()
        assert np.array_equal(final, final_local), "reference normal path != only_local path"
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            raw=raw.numpy(), feats=f.numpy(), q=sh.q, k1=k1, k2=k2, lam=lam,
            pids=torch.cat([qp, gp]).numpy(), cams=torch.cat([qc, gc]).numpy(),
            D=D if N <= 400 else np.zeros((0, 0), np.float32),
            final=final, eu=eu[:EU_ROWS], co=co[:EU_ROWS], cmc_rr=cmc_rr, map_rr=map_rr, cmc_eu=cmc_eu, map_eu=map_eu,
            cmc_eval=res[False][0], map_eval=res[False][1], cmc_eval_rr=res[True][0], map_eval_rr=res[True][1])
        print(name, final.shape, "mAP", map_eu, "->", map_rr)


if __name__ == "__main__":
    main()
