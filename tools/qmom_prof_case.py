"""One fused QMOM step (no S store) on 1e7 cells: the launch ncu captures (tools/qmom_prof_case.py)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
w = synthetic.workload("qmom3_ct", nc)
f = synthetic.qmom_fields(w, ncells=1_000_000)
rep = (nc + 999_999) // 1_000_000
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
m = T(np.tile(f["m"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc])
q = engine.QMOM(ctx, 3, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
d = torch.empty(nc, dtype=torch.float64, device=dev)
q.step(m, e, r, v, w.dt, d=d)
torch.cuda.synchronize()
print("ok", float(d.sum()))
