"""Timing of the implicit transport (fvm::ddt + fvm::div == S for all fields): python tools/implicit_time.py [n K reps courant tol solver scheme]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)
from openpbe_b200 import engine, synthetic
n = int(sys.argv[1]) if len(sys.argv) > 1 else 160
K = int(sys.argv[2]) if len(sys.argv) > 2 else 100
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
courant = float(sys.argv[4]) if len(sys.argv) > 4 else 0.4
tol = float(sys.argv[5]) if len(sys.argv) > 5 else 1e-5
solver = sys.argv[6] if len(sys.argv) > 6 else 'auto'
scheme = sys.argv[7] if len(sys.argv) > 7 else 'Gauss upwind'
ctx = engine.Context(0); dev = torch.device("cuda:0")
m = synthetic.block_mesh(n, n, n, velocity=(1.0, 0.5, 0.25))
mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
if scheme != "Gauss upwind": mesh.set_geometry(m["Sf"], m["C"], m["weights"])
nc = n ** 3
g = torch.Generator(device=dev); g.manual_seed(1)
psi = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
S = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
out = torch.empty_like(psi); phi = torch.from_numpy(m["phi"]).to(dev)
dt = courant / (n * 1.75)
for _ in range(2): _, perf = mesh.transport_implicit(psi, phi, dt, S=S, out=out, tolerance=tol, solver=solver, scheme=scheme)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); best = 1e9
for _ in range(reps):
    e0.record(); mesh.transport_implicit(psi, phi, dt, S=S, out=out, tolerance=tol, solver=solver, scheme=scheme); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
its = [p[2] for p in perf]
print(f"implicit {scheme} ({solver}) {n}^3 x {K}, Courant {courant}, tolerance {tol:g}: {best:.2f} ms, iterations {min(its)}..{max(its)}, "
      f"residuals {max(p[0] for p in perf):.2e} -> {max(p[1] for p in perf):.2e}, {nc*K/best*1e3:.3e} cell-field-updates/s, "
      f"{best/max(its):.2f} ms per iteration, checksum {float(out.sum()):.12e}")
