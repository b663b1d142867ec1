"""configs[0] (MOC 30 classes x 1e5 cells) step time under the kernel-selection env vars: python tools/config1_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic
ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
nc = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000
w = synthetic.workload("moc30_ct", nc)
f = synthetic.moc_fields(w)
n0, e, r, v = T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"])
moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
n = n0.clone()
g = torch.cuda.CUDAGraph()
for _ in range(5): moc.step(n, e, r, v, w.dt)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
reps = 200
e0.record()
for _ in range(reps): moc.step(n, e, r, v, w.dt)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"{moc.kernel()} {nc} cells: {ms*1e3:.1f} us per step (back to back, {reps} launches) -> {nc*w.N/ms*1e3:.3e} ccu/s")
