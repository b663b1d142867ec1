"""Smallest end-to-end case for compute-sanitizer: python tools/sanitize_case.py
MOC (30 classes, both kernels, source + fused step) and QMOM (order 3, step) on a few hundred cells."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic

ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for kern in ("tmem", "smem"):
    w = synthetic.workload("moc30_ct", 128 * 3 + 17)
    f = synthetic.moc_fields(w)
    moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
    moc.set_kernel(kern)
    n, e, r, v, vc, al = (T(f[k]) for k in ("n", "eps", "rho", "nu", "vcell", "alpha"))
    S = moc.source(n, e, r, v, w.dt)
    mom = torch.zeros(6, dtype=torch.float64, device=dev)
    d = torch.empty(n.shape[1], dtype=torch.float64, device=dev)
    moc.step(n, e, r, v, w.dt, d=d, vcell=vc, alpha=al, moments=mom)
    torch.cuda.synchronize(); print("moc", kern, float(S.abs().max()), mom.cpu().tolist()[:2])
wq = synthetic.workload("qmom3_ct", 128 * 2 + 5)
fq = synthetic.qmom_fields(wq)
q = engine.QMOM(ctx, 3, engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs))
m = T(fq["m"]); S = torch.empty_like(m); d = torch.empty(m.shape[1], dtype=torch.float64, device=dev)
q.step(m, T(fq["eps"]), T(fq["rho"]), T(fq["nu"]), wq.dt, S=S, d=d)
torch.cuda.synchronize(); print("qmom ok", float(d.mean()))
