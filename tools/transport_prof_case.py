"""Fixed launch sequence of the transport kernels for ncu: python tools/transport_prof_case.py [n]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic
n = int(sys.argv[1]) if len(sys.argv) > 1 else 160
ctx = engine.Context(0); dev = torch.device("cuda:0")
m = synthetic.block_mesh(n, n, n, velocity=(1.0, 0.5, 0.25))
mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
nc, K = n ** 3, 100
g = torch.Generator(device=dev); g.manual_seed(1)
psi = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
S = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
out = torch.empty_like(psi); phi = torch.from_numpy(m["phi"]).to(dev)
for _ in range(3):
    mesh.transport_explicit(psi, phi, 1e-4, S=S, out=out)
torch.cuda.synchronize()
print("transport ok", nc)
