"""Timing of the explicit transport update (row f2): python tools/transport_time.py [nx ny nz K]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import capi
capi.LIB_PATH = os.environ.get('PBE_LIB', capi.LIB_PATH)  # A/B builds of the library
from openpbe_b200 import engine, synthetic

nx, ny, nz, K = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (160, 160, 160, 100)
ctx = engine.Context(0); dev = torch.device("cuda:0")
t0 = time.time(); m = synthetic.block_mesh(nx, ny, nz, velocity=(1.0, 0.5, 0.25)); t1 = time.time()
mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"]); t2 = time.time()
nc = nx * ny * nz
print(f"mesh {nx}x{ny}x{nz}: {nc} cells, generator {t1-t0:.1f} s, pbe_mesh_create {t2-t1:.1f} s, {mesh.info()}")
g = torch.Generator(device=dev); g.manual_seed(1)
psi = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
S = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
out = torch.empty_like(psi)
phi = torch.from_numpy(m["phi"]).to(dev)
alg = nc * (24.0 * K + 8 + 6 * 8 + 3 * 8)
for chunk in (os.environ.get("CHUNKS", "5,10,25,50,100").split(",")):
    os.environ["PBE_TRANSPORT_CHUNK"] = chunk
    for _ in range(2): mesh.transport_explicit(psi, phi, 1e-4, S=S, out=out)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); best = 1e9
    for _ in range(5):
        e0.record(); mesh.transport_explicit(psi, phi, 1e-4, S=S, out=out); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"chunk {chunk}: {best:.3f} ms, {nc*K/best*1e3:.3e} cell-field-updates/s, algorithmic {alg/best/1e6:.0f} GB/s")
