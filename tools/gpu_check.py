"""Quick GPU sanity run: parity vs the oracle on small cases, then timings. Not a test."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic
from oracle import cpu_oracle as orc

ctx = engine.Context(0)
print(ctx.info())
print("fp64 peak TF/s, ms:", ctx.fp64_peak_tflops())
dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

def moc_case(name, ncells):
    w = synthetic.workload(name, ncells)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    ospec = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    S = moc.source(T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"]), w.dt).cpu().numpy()
    So, scale = orc.moc_source(ospec, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt, nthreads=8)
    err = np.abs(S - So) / (scale + 1e-300)
    print(f"{name} N={w.N} cells={ncells}: max |dS|/scale = {err.max():.3e}  nan={np.isnan(S).sum()} limited={(So == -w.relax*f['n']/w.dt).mean():.3f}")
    return moc, w

for nm in ("moc30_ct", "moc100_ct", "moc100_const", "replay_moc"):
    moc_case(nm, 1000)

def qmom_case(name, ncells):
    w = synthetic.workload(name, ncells)
    f = synthetic.qmom_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    ospec = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    S = q.source(T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()
    So, scale = orc.qmom_source(ospec, w.N, f["m"], f["eps"], f["rho"], f["nu"], nthreads=8)
    ok = np.isfinite(So)
    err = np.abs(S - So)[ok] / (scale[ok] + 1e-300)
    print(f"{name}: max |dS|/scale = {err.max():.3e} median {np.median(err):.2e} nan match={np.array_equal(np.isnan(S), np.isnan(So))} nans={np.isnan(So).sum()}")
for nm in ("qmom3_ct", "qmom3_const"):
    qmom_case(nm, 20000)

# timings
def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    best = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best

for nm, nc in (("moc100_ct", 1_600_000), ("moc100_const", 1_600_000), ("moc30_ct", 1_000_000)):
    w = synthetic.workload(nm, nc)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    n, e, r, v = T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"])
    S = torch.empty_like(n)
    ms = timeit(lambda: moc.source(n, e, r, v, w.dt, out=S))
    print(f"{nm} cells={nc}: {ms:.3f} ms -> {nc*w.N/ms*1e3:.3e} cell-class-updates/s, {nc*(2*w.N+3)*8/ms/1e6:.1f} GB/s")
w = synthetic.workload("qmom3_ct", 10_000_000)
f = synthetic.qmom_fields(w)
spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
q = engine.QMOM(ctx, w.N, spec)
m, e, r, v = T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])
S = torch.empty_like(m)
ms = timeit(lambda: q.source(m, e, r, v, out=S))
print(f"qmom3_ct cells=1e7: {ms:.3f} ms -> {1e7/ms*1e3:.3e} cell-updates/s, {1e7*120/ms/1e6:.1f} GB/s")
