"""Symmetric-schedule ablation in CYCLES (clock probe), interleaved with the baseline to cancel thermal drift."""
import importlib, os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
engine = importlib.import_module("multimodal-reid_b200.engine")
N, d = int(os.environ.get("N", 19889)), 512
eng = engine.Engine(torch.device("cuda", 0))
f = eng.prepare(torch.randn(N, d, device="cuda"), True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def cycles(knob, sym=True):
    os.environ["REID_TC_DBG"] = str(knob | 64)
    eng.symmetric = sym
    out = (ctypes.c_int64 * 4)()
    cs = []
    for i in range(6):
        flush.fill_(1)
        eng.dist(f, f, 0, want_rowmax=True, want_chunk_min=True)
        torch.cuda.synchronize()
        eng.lib.reid_dist_clock_probe(out)
        cs.append(out[2] - out[0])
    return sorted(cs[2:])[len(cs[2:]) // 2] / 1e3
for knob in [int(k) for k in os.environ.get("KNOBS", "0,1,2,3,4,8,16,32,34,63").split(",")]:
    print(f"dbg={knob:3d}  sym {cycles(knob):7.0f} kcyc   full {cycles(knob, False):7.0f} kcyc", flush=True)
