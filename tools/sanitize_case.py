"""Smallest end-to-end case for compute-sanitizer: python tools/sanitize_case.py
MOC (30 classes, both kernels, source + fused step) and QMOM (order 3, step) on a few hundred cells."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openpbe_b200 import engine, synthetic

ctx = engine.Context(0); dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
for kern in ("tmem", "smem"):
    w = synthetic.workload("moc30_ct", 128 * 3 + 17)
    f = synthetic.moc_fields(w)
    moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
    moc.set_kernel(kern)
    n, e, r, v, vc, al = (T(f[k]) for k in ("n", "eps", "rho", "nu", "vcell", "alpha"))
    S = moc.source(n, e, r, v, w.dt)
    mom = torch.zeros(6, dtype=torch.float64, device=dev)
    d = torch.empty(n.shape[1], dtype=torch.float64, device=dev)
    moc.step(n, e, r, v, w.dt, d=d, vcell=vc, alpha=al, moments=mom)
    torch.cuda.synchronize(); print("moc", kern, float(S.abs().max()), mom.cpu().tolist()[:2])
wq = synthetic.workload("qmom3_ct", 128 * 2 + 5)
fq = synthetic.qmom_fields(wq)
q = engine.QMOM(ctx, 3, engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs))
m = T(fq["m"]); S = torch.empty_like(m); d = torch.empty(m.shape[1], dtype=torch.float64, device=dev)
q.step(m, T(fq["eps"]), T(fq["rho"]), T(fq["nu"]), wq.dt, S=S, d=d)
torch.cuda.synchronize(); print("qmom ok", float(d.mean()))

# ---- round 2: fixed-pivot kernels (tensor-memory and shared-memory accumulators), MULES (fused and per-field forms),
# resident correct(), the allocation-free min / max
import ctypes as C
from openpbe_b200 import capi
for pk in ("tmem", "smem"):
    os.environ["PBE_MOC_PIVOT_KERNEL"] = pk
    w = synthetic.workload("moc30_ct", 128 * 2 + 9)
    f = synthetic.moc_fields(w)
    x = w.xi1 * 1.3 ** np.arange(w.N)
    moc = engine.MOC(ctx, w.N, 0.0, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax, pivots=x)
    n, e, r, v = (T(f[k]) for k in ("n", "eps", "rho", "nu"))
    S = moc.source(n, e, r, v, w.dt)
    mom = torch.zeros(3, dtype=torch.float64, device=dev)
    moc.step(n, e, r, v, w.dt, moments=mom)
    torch.cuda.synchronize(); print("pivot", pk, moc.kernel(), float(S.abs().max()))
os.environ.pop("PBE_MOC_PIVOT_KERNEL")
m = synthetic.block_mesh(9, 7, 5, velocity=(1.0, -0.5, 0.25))
mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
mesh.set_geometry(m["Sf"], m["C"], m["weights"])
K, nc = 5, 9 * 7 * 5
g = torch.Generator(device=dev); g.manual_seed(3)
psi = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
Ssrc = torch.rand(K, nc, dtype=torch.float64, device=dev, generator=g)
pmax = torch.full((K,), 0.9, dtype=torch.float64, device=dev)
phi = T(m["phi"])
for path in ("1", "0"):
    os.environ["PBE_MULES_PATH"] = path
    out = mesh.transport_mules(psi, phi, 0.01, pmax, S=Ssrc)
    torch.cuda.synchronize(); print("mules path", path, float(out.sum()))
os.environ.pop("PBE_MULES_PATH")
out = mesh.transport_explicit(psi, phi, 0.01, S=Ssrc)
w = synthetic.workload("moc30_ct", nc)
f = synthetic.moc_fields(w)
moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
for scheme in ("upwind", "vanLeer"):
    st = engine.MOCState(moc, nc, vcell=m["vcell"], n0=f["n"], mesh=mesh, scheme=scheme)
    d, diag = st.correct(f["eps"], f["rho"], f["nu"], f["alpha"], w.dt, phi=m["phi"], minmax=True, n_moments=3)
    torch.cuda.synchronize(); print("resident", scheme, float(d.mean()), diag["cont_err_max"])
    st.close()
nn = T(f["n"]); lo = torch.empty(w.N, dtype=torch.float64, device=dev); hi = torch.empty_like(lo)
rc = capi.lib().pbe_moc_class_minmax_noalloc(moc._h, nc, nn.stride(0), C.c_void_p(nn.data_ptr()), C.c_void_p(lo.data_ptr()),
                                             C.c_void_p(hi.data_ptr()), None)
torch.cuda.synchronize(); print("noalloc minmax rc", rc, float(lo.min()), float(hi.max()))
