"""A/B timing of the two MOC device kernels on one GPU (not a bench: quick CUDA-event timings)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)  # A/B builds of the library
from openpbe_b200 import engine, synthetic

ctx = engine.Context(0)
dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    best = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best

cases = [("moc100_ct", 148 * 256 * 40), ("moc100_const", 148 * 256 * 40), ("replay_moc", 148 * 256 * 40), ("moc30_ct", 148 * 256 * 80)]
if len(sys.argv) > 1:
    cases = [(sys.argv[1], int(float(sys.argv[2])))]
for nm, nc in cases:
    w = synthetic.workload(nm, nc)
    f = synthetic.moc_fields(w, ncells=min(nc, 200_000))
    rep = (nc + f["n"].shape[1] - 1) // f["n"].shape[1]
    n = T(np.tile(f["n"], (1, rep))[:, :nc]); e = T(np.tile(f["eps"], rep)[:nc]); r = T(np.tile(f["rho"], rep)[:nc]); v = T(np.tile(f["nu"], rep)[:nc])
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S = torch.empty_like(n)
    out = {}
    for k in ("smem", "tmem"):
        moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
        try:
            moc.set_kernel(k)
        except Exception as ex:
            print(nm, k, "unavailable"); continue
        ms = timeit(lambda: moc.source(n, e, r, v, w.dt, out=S))
        out[k] = S.clone()
        print(f"{nm} N={w.N} cells={nc} {k} {moc.kernel()}: {ms:.3f} ms -> {nc*w.N/ms*1e3:.3e} ccu/s")
    if len(out) == 2:
        d = (out["smem"] - out["tmem"]).abs().max().item(); m = out["smem"].abs().max().item()
        print(f"   max |S_smem - S_tmem| = {d:.3e} (max |S| {m:.3e})")
