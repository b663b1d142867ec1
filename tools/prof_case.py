"""Small fixed launch sequence for ncu: `python tools/prof_case.py moc|qmom|all [cells]`.

Not a test and not a bench: no number printed here is a bench value.  The launch list / full
captures taken from this script are summarised under profiles/.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from openpbe_b200 import engine, synthetic  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
ctx = engine.Context(0)
dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731

if what in ("moc", "all"):
    nc = int(float(sys.argv[2])) if len(sys.argv) > 2 else 148 * 128 * 16
    w = synthetic.workload("moc100_ct", nc)
    f = synthetic.moc_fields(w)
    moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs),
                     w.relax)
    n, e, r, v, vc = T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"]), T(f["vcell"])
    mom = torch.zeros(6, dtype=torch.float64, device=dev)
    for _ in range(3):
        moc.step(n, e, r, v, w.dt, vcell=vc, moments=mom)
    torch.cuda.synchronize()
    print("moc ok", nc)

if what in ("qmom", "all"):
    nc = int(float(sys.argv[2])) if len(sys.argv) > 2 and what == "qmom" else 4_000_000
    w = synthetic.workload("qmom3_ct", nc)
    f = synthetic.qmom_fields(w)
    q = engine.QMOM(ctx, w.N, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    m, e, r, v = T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])
    S = torch.empty_like(m)
    d = torch.empty(nc, dtype=torch.float64, device=dev)
    for _ in range(3):
        q.step(m, e, r, v, w.dt, S=S, d=d)
    torch.cuda.synchronize()
    print("qmom ok", nc)
