"""Quick timing of the fused MOC step at bench shape on a smaller field: python tools/moc_step_time.py [cells]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)  # A/B builds of the library
import numpy as np, torch
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 148 * 256 * 160
w = synthetic.workload("moc100_ct", nc)
f = synthetic.moc_fields(w, ncells=200_000); rep = (nc + 199_999) // 200_000
n = T(np.tile(f["n"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc]); vc = T(np.tile(f["vcell"], rep)[:nc])
moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
mom = torch.zeros(6, dtype=torch.float64, device=dev)
for _ in range(2): moc.step(n, e, r, v, w.dt, vcell=vc, moments=mom)
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
best = 1e9
for _ in range(4):
    e0.record(); moc.step(n, e, r, v, w.dt, vcell=vc, moments=mom); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"moc100_ct step {nc} cells: {best:.3f} ms -> {nc*w.N/best*1e3:.4e} ccu/s")
