"""Generate tests/golden/*.npz: small seeded inputs with the oracle's outputs.

The reference is C++ against OpenFOAM / Eigen3 / GSL and cannot be built or imported here
(DESIGN.md), so these vectors come from oracle/pbe_oracle.c (C restatement, pinned on the
reference's unit-test goldens in reference_unit_tests.json) and are cross-checked, before they
are written, against the independent NumPy restatement oracle/numpy_restatement.py.
Run from the repo root:  python tests/golden/make_fixtures.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cpu_oracle as orc  # noqa: E402
from oracle import numpy_restatement as npo  # noqa: E402
from openpbe_b200 import synthetic  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def moc_fixture(name, wl, ncells):
    w = synthetic.workload(wl, ncells)
    f = synthetic.moc_fields(w)
    spec = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S, scale, coal, br = orc.moc_source(spec, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"],
                                        w.dt, want_parts=True)
    Sn = npo.moc_source(dict(coalescence=w.coalescence, coal_c=w.coal_coeffs, breakup=w.breakup,
                             br_c=w.br_coeffs), w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
    assert np.all(np.abs(S - Sn) <= 1e-12 * scale + 1e-300), name
    n1 = f["n"].copy()
    orc.moc_update(n1, S, w.dt)
    d1 = orc.moc_d(n1, w.xi1)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), workload=wl, N=w.N, xi1=w.xi1, dt=w.dt,
                        relax=w.relax, n=f["n"], eps=f["eps"], rho=f["rho"], nu=f["nu"], S=S, scale=scale,
                        n_after=n1, d_after=d1)
    print(name, S.shape, "ok")


def qmom_fixture(name, wl, ncells):
    w = synthetic.workload(wl, ncells)
    f = synthetic.qmom_fields(w)
    spec = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S, scale = orc.qmom_source(spec, w.N, f["m"], f["eps"], f["rho"], f["nu"])
    nspec = dict(coalescence=w.coalescence, coal_c=w.coal_coeffs, breakup=w.breakup, br_c=w.br_coeffs)
    ncheck = min(ncells, 200)  # pure-Python per-cell loop: a sample is enough
    Sn = np.full_like(S, np.nan)
    with np.errstate(all="ignore"):
        for c in range(ncheck):
            Sn[:, c] = npo.qmom_source_cell(nspec, w.N, f["m"][:, c], f["eps"][c], f["rho"][c], f["nu"][c])
    well = (f["nodes"][1] / f["nodes"][0] > 1.001)
    ok = np.isfinite(S) & np.isfinite(Sn) & well[None, :]
    assert np.all(np.abs(S - Sn)[ok] <= 1e-7 * scale[ok] + 1e-300), name
    x = np.empty((w.N, ncells))
    wts = np.empty((w.N, ncells))
    for c in range(ncells):
        x[:, c], wts[:, c] = orc.wheeler(f["m"][:, c])
    m1 = f["m"].copy()
    S1 = S.copy()
    d1 = orc.qmom_update(m1, S1, w.dt)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), workload=wl, order=w.N, dt=w.dt, m=f["m"],
                        eps=f["eps"], rho=f["rho"], nu=f["nu"], S=S, scale=scale, x=x, w=wts,
                        well=well, m_after=m1, S_scrubbed=S1, d_after=d1)
    print(name, S.shape, "ok")


if __name__ == "__main__":
    orc.build()
    moc_fixture("moc30_ct_257", "moc30_ct", 257)
    moc_fixture("moc100_ct_64", "moc100_ct", 64)
    moc_fixture("moc50_c01_96", "replay_moc", 96)
    moc_fixture("moc100_const_40", "moc100_const", 40)
    qmom_fixture("qmom3_ct_1000", "qmom3_ct", 1000)
    qmom_fixture("qmom3_const_500", "qmom3_const", 500)
