"""Phase split of the MOC tensor-memory kernel (needs a library built with -DPBE_PHASE_TIMING,
passed as PBE_LIB): cycles of thread 0 of every CTA in staging / pair sweep / epilogue."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)
import numpy as np, torch
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = 148 * 256 * 40
w = synthetic.workload("moc100_ct", nc)
f = synthetic.moc_fields(w, ncells=200_000); rep = (nc + 199_999) // 200_000
n = T(np.tile(f["n"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc]); vc = T(np.tile(f["vcell"], rep)[:nc])
moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
mom = torch.zeros(8, dtype=torch.float64, device=dev)
moc.step(n, e, r, v, w.dt, vcell=vc, moments=mom); mom.zero_()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record(); moc.step(n, e, r, v, w.dt, vcell=vc, moments=mom); e1.record(); torch.cuda.synchronize()
m = mom.cpu().numpy(); ph = m[5:8]; nb = nc / 128
print(f"{e0.elapsed_time(e1):.3f} ms; cycles per batch: staging {ph[0]/nb:.0f}, pairs {ph[1]/nb:.0f}, epilogue {ph[2]/nb:.0f}; shares {ph/ph.sum()}")
