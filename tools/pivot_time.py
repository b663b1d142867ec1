"""Timing of the fixed-pivot MOC step on a geometric grid: python tools/pivot_time.py [cells] [ratio]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 148 * 128 * 16
ratio = float(sys.argv[2]) if len(sys.argv) > 2 else 1.08
w = synthetic.workload("moc100_ct", nc)
f = synthetic.moc_fields(w, ncells=min(nc, 200_000)); rep = (nc + f["n"].shape[1] - 1) // f["n"].shape[1]
x = w.xi1 * ratio ** np.arange(w.N)
n = T(np.tile(f["n"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc])
spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
for name, moc in (("pivot", engine.MOC(ctx, w.N, 0.0, spec, w.relax, pivots=x)), ("uniform", engine.MOC(ctx, w.N, w.xi1, spec, w.relax))):
    mom = torch.zeros(6, dtype=torch.float64, device=dev)
    for _ in range(2): moc.step(n.clone(), e, r, v, w.dt, moments=mom)
    nn = n.clone(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); best = 1e9
    for _ in range(3):
        e0.record(); moc.step(nn, e, r, v, w.dt, moments=mom); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"{name} grid ({moc.kernel()}), N {w.N}, ratio {ratio}: {nc} cells {best:.3f} ms -> {nc*w.N/best*1e3:.4e} cell-class-updates/s")
