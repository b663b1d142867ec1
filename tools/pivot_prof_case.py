"""Fixed launch sequence of the fixed-pivot kernel for ncu: python tools/pivot_prof_case.py [cells] [ratio]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 148 * 128 * 8
ratio = float(sys.argv[2]) if len(sys.argv) > 2 else 1.08
w = synthetic.workload("moc100_ct", nc)
f = synthetic.moc_fields(w, ncells=min(nc, 200_000)); rep = (nc + f["n"].shape[1] - 1) // f["n"].shape[1]
x = w.xi1 * ratio ** np.arange(w.N)
n = T(np.tile(f["n"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc])
moc = engine.MOC(ctx, w.N, 0.0, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax, pivots=x)
mom = torch.zeros(6, dtype=torch.float64, device=dev)
for _ in range(3): moc.step(n, e, r, v, w.dt, moments=mom)
torch.cuda.synchronize(); print("pivot ok", nc)
