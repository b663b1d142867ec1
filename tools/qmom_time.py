"""Quick CUDA-event timing of the QMOM kernels (not a bench): python tools/qmom_time.py [cells]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)  # A/B builds of the library
from openpbe_b200 import engine, synthetic

ctx = engine.Context(0)
dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for nm, order in (("qmom3_ct", 3), ("qmom3_const", 3), ("qmom3_ct", 2), ("qmom3_ct", 4)):
    w = synthetic.workload(nm, nc)
    w.N = order
    f = synthetic.qmom_fields(w, ncells=min(nc, 500_000))
    rep = (nc + 499_999) // 500_000
    m0 = T(np.tile(f["m"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc])
    q = engine.QMOM(ctx, order, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    S = torch.empty_like(m0); d = torch.empty(nc, dtype=torch.float64, device=dev)
    best = {}
    for what in ("source", "step"):
        t = 1e9
        for _ in range(5):
            m = m0.clone()
            if not os.environ.get('NOFLUSH'): flush.zero_()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            if what == "source":
                q.source(m, e, r, v, out=S)
            else:
                q.step(m, e, r, v, w.dt, S=S, d=d)
            e1.record(); torch.cuda.synchronize()
            t = min(t, e0.elapsed_time(e1))
        best[what] = t
    print(f"{nm} order {order} cells {nc}: source {best['source']:.3f} ms, step {best['step']:.3f} ms -> {nc/best['step']*1e3:.3e} cell-updates/s")
