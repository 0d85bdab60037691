import importlib, sys, torch, numpy as np
sys.path.insert(0, "/root/repo")
engine = importlib.import_module("multimodal-reid_b200.engine")
eng = engine.Engine()
torch.manual_seed(0)
n, d = 2048, 512
X = torch.randn(n, d, device="cuda")
X[1::2] = X[0::2] + 0.3 * torch.randn(n // 2, d, device="cuda")      # correlated pairs -> large dots
f = eng.prepare(X, True)
e = 14 - int(np.floor(np.log2(f.amax.cpu().numpy().view(np.float32)[0])))
hi = f.hi.double() * 2.0 ** -e
lo = f.lo.double() * 2.0 ** -e
x = f.x32.double()
sq = f.sq.double()
def run(fe):
    D, _ = eng.dist(fe, fe)
    return ((sq[:, None] + sq[None, :]) - D[:, :n].double()) / 2.0     # recovered dot (approx; sq rounding adds ~1e-7)
# 1) full split
dot = run(f)
ref = x @ x.T
err = dot - ref
big = ref.abs() > 0.5
print("full : max|err| %.3e  mean err (dot>0.5) %.3e  (dot<-0.0) n/a  rms %.3e" % (err.abs().max(), err[ref > 0.5].mean(), err.pow(2).mean().sqrt()))
# 2) hi only: zero lo
f2 = engine.Features(f.x32, f.hi, torch.zeros_like(f.lo), f.sq, f.nrm, f.amax, f.d)
dot2 = run(f2)
ref2 = hi @ hi.T
err2 = dot2 - ref2
print("hi.hi: max|err| %.3e  mean err (dot>0.5) %.3e  rms %.3e" % (err2.abs().max(), err2[ref2 > 0.5].mean(), err2.pow(2).mean().sqrt()))
# negative dots: flip sign of odd rows
f3 = engine.Features(f.x32, torch.where((torch.arange(n, device="cuda") % 2 == 1)[:, None], -f.hi, f.hi), torch.zeros_like(f.lo), f.sq, f.nrm, f.amax, f.d)
hi3 = f3.hi.double() * 2.0 ** -e
dot3 = run(f3)
ref3 = hi3 @ hi3.T
err3 = dot3 - ref3
print("neg  : mean err (dot<-0.5) %.3e   mean err (dot>0.5) %.3e" % (err3[ref3 < -0.5].mean(), err3[ref3 > 0.5].mean()))
# error of the epilogue alone (sq + sq - 2 dot in fp32) is ~1e-7: measure on D directly
D, _ = eng.dist(f, f)
refD = (sq[:, None] + sq[None, :]) - 2 * ref
eD = D[:, :n].double() - refD
print("D    : max|err| %.3e  mean (dot>0.5) %.3e" % (eD.abs().max(), eD[ref > 0.5].mean()))
# SIMT engine for comparison
D2, _ = eng.dist(f, f, engine=engine.ENGINES["simt"])
eD2 = D2[:, :n].double() - refD
print("simt : max|err| %.3e  mean (dot>0.5) %.3e" % (eD2.abs().max(), eD2[ref > 0.5].mean()))
