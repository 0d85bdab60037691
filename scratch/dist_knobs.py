"""Times the symmetric N x N contraction alone (Duke shape) under the REID_TC_DBG ablation knobs."""
import importlib, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
engine = importlib.import_module("multimodal-reid_b200.engine")
N, d = int(os.environ.get("N", 19889)), 512
eng = engine.Engine(torch.device("cuda", 0))
x = torch.randn(N, d, device="cuda")
f = eng.prepare(x, True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def run(sym=True):
    eng.symmetric = sym
    ts = []
    for i in range(8):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.dist(f, f, 0, want_rowmax=True, want_chunk_min=True); b.record()
        torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts[3:])[len(ts[3:]) // 2]
import ctypes
def probe():
    out = (ctypes.c_int64 * 4)()
    eng.lib.reid_dist_clock_probe(out)
    cyc, ns = out[2] - out[0], out[3] - out[1]
    return cyc / max(ns, 1) * 1e3, ns / 1e6
for knob in [int(k) for k in os.environ.get("KNOBS", "0,4,16,63").split(",")]:
    os.environ["REID_TC_DBG"] = str(knob | 64)
    a = run(True); pa = probe(); b = run(False); pb = probe()
    print(f"dbg={knob:3d}  sym {a:.4f} ms ({pa[0]:.0f} MHz, blk0 {pa[1]:.3f} ms)   full {b:.4f} ms ({pb[0]:.0f} MHz, blk0 {pb[1]:.3f} ms)", flush=True)
