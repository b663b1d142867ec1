"""Timing of the explicit upwind transport update: python tools/transport_time.py [n K reps]  (PBE_TRANSPORT_KERNEL, PBE_TRANSPORT_CHUNK)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)
from openpbe_b200 import engine, synthetic
n = int(sys.argv[1]) if len(sys.argv) > 1 else 160
K = int(sys.argv[2]) if len(sys.argv) > 2 else 100
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
ctx = engine.Context(0); dev = torch.device("cuda:0")
m = synthetic.block_mesh(n, n, n, velocity=(1.0, 0.5, 0.25))
mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
nc = n ** 3
g = torch.Generator(device=dev); g.manual_seed(1)
psi = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
S = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
out = torch.empty_like(psi); phi = torch.from_numpy(m["phi"]).to(dev)
dt = 0.4 / (n * 1.75)
for _ in range(3): mesh.transport_explicit(psi, phi, dt, S=S, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(reps): mesh.transport_explicit(psi, phi, dt, S=S, out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"explicit upwind {n}^3 x {K} (kernel {os.environ.get('PBE_TRANSPORT_KERNEL', 'default')}, chunk {os.environ.get('PBE_TRANSPORT_CHUNK', 'default')}): "
      f"{ms:.3f} ms, {24.0 * nc * K / ms / 1e6:.0f} GB/s, checksum {float(out.sum()):.12e}")
