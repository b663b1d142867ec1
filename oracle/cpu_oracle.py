"""ctypes loader for oracle/_build/libpbe_oracle.so (the C restatement of the reference).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never by the product package `openpbe_b200`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libpbe_oracle.so")

COAL_NONE, COAL_CONSTANT, COAL_CT, COAL_TABLE = 0, 1, 2, 3
BR_NONE, BR_BINARY, BR_CT, BR_TABLE = 0, 1, 2, 3
DAUGHTER_UNIFORM_BINARY = 0

_dp = C.POINTER(C.c_double)


class KernelSpec(C.Structure):
    _fields_ = [
        ("coalescence", C.c_int),
        ("coal_c", C.c_double * 4),
        ("breakup", C.c_int),
        ("br_c", C.c_double * 4),
        ("daughter", C.c_int),
        ("coal_div_vcell", C.c_int),
        ("coal_table", _dp),
        ("br_table", _dp),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "pbe_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(src),
                                                 os.path.getmtime(os.path.join(_HERE, "pbe_oracle.h")))):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    L = C.CDLL(_LIB_PATH)
    d, i, i64 = C.c_double, C.c_int, C.c_int64
    L.orc_coal_ct.restype = d
    L.orc_coal_ct.argtypes = [d] * 10 + [i]
    L.orc_coal_constant.restype = d
    L.orc_coal_constant.argtypes = [d]
    L.orc_coal_none.restype = d
    L.orc_break_ct.restype = d
    L.orc_break_ct.argtypes = [d] * 7
    L.orc_break_binary.restype = d
    L.orc_break_binary.argtypes = [d]
    L.orc_break_none.restype = d
    L.orc_break_none.argtypes = [d]
    L.orc_daughter_beta.restype = d
    L.orc_daughter_beta.argtypes = [d, d]
    L.orc_daughter_moment.restype = d
    L.orc_daughter_moment.argtypes = [d, i]
    L.orc_moc_source.restype = i
    L.orc_moc_source.argtypes = [C.POINTER(KernelSpec), i, d, d, i64, i64, _dp, _dp, _dp, _dp, _dp, d,
                                 _dp, _dp, _dp, _dp, i]
    L.orc_moc_update.restype = None
    L.orc_moc_update.argtypes = [i, i64, i64, _dp, _dp, d]
    L.orc_moc_d.restype = None
    L.orc_moc_d.argtypes = [i, d, i64, i64, _dp, _dp]
    L.orc_moc_continuity.restype = d
    L.orc_moc_continuity.argtypes = [i, d, i64, i64, _dp, _dp, _dp, _dp]
    L.orc_wheeler.restype = None
    L.orc_wheeler.argtypes = [i, _dp, _dp, _dp]
    L.orc_tridiag_eig_firstrow.restype = i
    L.orc_tridiag_eig_firstrow.argtypes = [i, _dp, _dp, _dp]
    L.orc_qmom_source.restype = i
    L.orc_qmom_source.argtypes = [C.POINTER(KernelSpec), i, i64, i64, _dp, _dp, _dp, _dp, _dp, _dp,
                                  _dp, i, i]
    L.orc_qmom_update.restype = None
    L.orc_qmom_update.argtypes = [i, i64, i64, _dp, _dp, d, _dp]
    L.orc_integrate_exp_neg.restype = d
    L.orc_integrate_exp_neg.argtypes = [d]
    L.orc_max_threads.restype = i
    ip = C.POINTER(C.c_int32)
    L.orc_transport_upwind_explicit.restype = None
    L.orc_transport_upwind_explicit.argtypes = [i64, i64, ip, ip, i64, ip, _dp, i, i64, _dp, _dp, _dp,
                                                i64, _dp, d, _dp]
    L.orc_transport_upwind_explicit_coupled.restype = None
    L.orc_transport_upwind_explicit_coupled.argtypes = [i64, i64, ip, ip, i64, ip, _dp, i, i64, _dp, _dp,
                                                        _dp, i64, i64, i64, _dp, i64, _dp, d, _dp]
    _lib = L
    return L


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def make_spec(coalescence="noCoalescence", coal_coeffs=(), breakup="noBreakup", br_coeffs=(),
              coal_div_vcell=False, coal_table=None, br_table=None) -> KernelSpec:
    """Kernel selection by the reference's TypeNames (coalescenceKernel.C:46-71 etc.)."""
    cmap = {"noCoalescence": COAL_NONE, "constantCoalescence": COAL_CONSTANT,
            "CoulaloglouTavlaridesC": COAL_CT, "table": COAL_TABLE}
    bmap = {"noBreakup": BR_NONE, "binaryBreakup": BR_BINARY, "CoulaloglouTavlarides": BR_CT,
            "table": BR_TABLE}
    s = KernelSpec()
    s.coalescence = cmap[coalescence]
    s.breakup = bmap[breakup]
    for k, v in enumerate(coal_coeffs):
        s.coal_c[k] = float(v)
    for k, v in enumerate(br_coeffs):
        s.br_c[k] = float(v)
    s.daughter = DAUGHTER_UNIFORM_BINARY
    s.coal_div_vcell = int(bool(coal_div_vcell))
    s._keep = (coal_table, br_table)
    s.coal_table = _p(coal_table)
    s.br_table = _p(br_table)
    return s


def moc_source(spec, N, xi1, relax, n, eps, rho, nu, dt, vcell=None, nthreads=1, want_parts=False):
    """n: float64 [N, ncells] class-major.  Returns (S, scale[, coal, br])."""
    n = np.ascontiguousarray(n, dtype=np.float64)
    assert n.shape[0] == N
    ncells = n.shape[1]
    S = np.empty_like(n)
    scale = np.empty_like(n)
    coal = np.empty_like(n) if want_parts else None
    br = np.empty_like(n) if want_parts else None
    rc = lib().orc_moc_source(C.byref(spec), N, xi1, relax, ncells, ncells, _p(n), _p(eps), _p(rho),
                              _p(nu), _p(vcell), dt, _p(S), _p(scale), _p(coal), _p(br), nthreads)
    if rc != 0:
        raise RuntimeError(f"orc_moc_source failed: {rc}")
    return (S, scale, coal, br) if want_parts else (S, scale)


def moc_update(n, S, dt):
    N, ncells = n.shape
    lib().orc_moc_update(N, ncells, ncells, _p(n), _p(S), dt)


def moc_d(n, xi1):
    N, ncells = n.shape
    d = np.empty(ncells)
    lib().orc_moc_d(N, xi1, ncells, ncells, _p(n), _p(d))
    return d


def moc_continuity(n, xi1, alpha, vcell):
    N, ncells = n.shape
    err = np.empty(ncells)
    mx = lib().orc_moc_continuity(N, xi1, ncells, ncells, _p(n), _p(alpha), _p(vcell), _p(err))
    return mx, err


def wheeler(m):
    m = np.ascontiguousarray(m, dtype=np.float64)
    order = m.size // 2
    x = np.empty(order)
    w = np.empty(order)
    lib().orc_wheeler(order, _p(m), _p(x), _p(w))
    return x, w


def tridiag_eig_firstrow(d, e):
    n = len(d)
    dd = np.array(d, dtype=np.float64)
    ee = np.zeros(n)
    ee[: n - 1] = e
    z = np.empty(n)
    rc = lib().orc_tridiag_eig_firstrow(n, _p(dd), _p(ee), _p(z))
    return rc, dd, z


def qmom_source(spec, order, m, eps, rho, nu, vcell=None, reference_shaped=False, nthreads=1):
    m = np.ascontiguousarray(m, dtype=np.float64)
    assert m.shape[0] == 2 * order
    ncells = m.shape[1]
    S = np.empty_like(m)
    scale = np.empty_like(m)
    rc = lib().orc_qmom_source(C.byref(spec), order, ncells, ncells, _p(m), _p(eps), _p(rho), _p(nu),
                               _p(vcell), _p(S), _p(scale), int(reference_shaped), nthreads)
    if rc != 0:
        raise RuntimeError(f"orc_qmom_source failed: {rc}")
    return S, scale


def qmom_update(m, S, dt, want_d=True):
    nm, ncells = m.shape
    d = np.empty(ncells) if want_d else None
    lib().orc_qmom_update(nm // 2, ncells, ncells, _p(m), _p(S), dt, _p(d))
    return d


def transport_upwind_explicit(owner, neighbour, boundary_owner, vcell, psi, phi, dt, psi_b=None, S=None,
                              coupled=None, psi_halo=None):
    """MULES::explicitSolve with the upwind flux on OpenFOAM face addressing (see pbe_oracle.h;
    parity unpinned: OpenFOAM itself is absent).  psi [K][ncells] -> new psi.  coupled = (first, count)
    marks processor-patch boundary faces whose across-the-face cell values are psi_halo [K][count]."""
    ip = C.POINTER(C.c_int32)
    own, nei, bown = (np.ascontiguousarray(a, dtype=np.int32) for a in (owner, neighbour, boundary_owner))
    psi = np.ascontiguousarray(psi, dtype=np.float64)
    K, ncells = psi.shape
    out = np.empty_like(psi)
    if psi_b is not None:
        psi_b = np.ascontiguousarray(psi_b, dtype=np.float64)
    c0, nc = coupled if coupled is not None else (0, 0)
    if nc:
        psi_halo = np.ascontiguousarray(psi_halo, dtype=np.float64)
        assert psi_halo.shape == (K, nc)
    lib().orc_transport_upwind_explicit_coupled(
        ncells, own.size, own.ctypes.data_as(ip), nei.ctypes.data_as(ip), bown.size,
        bown.ctypes.data_as(ip), _p(np.ascontiguousarray(vcell, dtype=np.float64)), K, ncells, _p(psi),
        _p(np.ascontiguousarray(phi, dtype=np.float64)), _p(psi_b), bown.size, c0, nc,
        _p(psi_halo) if nc else None, nc, _p(S), float(dt), _p(out))
    return out


# ---- MULES-limited transport (oracle/pbe_oracle_mules.c; parity unpinned: OpenFOAM restated) ----------
class OrcMesh(C.Structure):
    _fields_ = [("ncells", C.c_int64), ("nint", C.c_int64), ("nbnd", C.c_int64), ("owner", C.c_void_p),
                ("neighbour", C.c_void_p), ("bown", C.c_void_p), ("V", C.c_void_p), ("Sf", C.c_void_p),
                ("C", C.c_void_p), ("weights", C.c_void_p)]


DIV_SCHEMES = {"upwind": 0, "Gauss upwind": 0, "vanLeer": 1, "Gauss vanLeer": 1}


def _orc_mesh(mesh: dict):
    keep = {k: np.ascontiguousarray(mesh[k], dtype=np.int32) for k in ("owner", "neighbour", "boundary_owner")}
    keep["V"] = np.ascontiguousarray(mesh["vcell"], dtype=np.float64)
    for k in ("Sf", "C", "weights"):
        keep[k] = np.ascontiguousarray(mesh[k], dtype=np.float64) if mesh.get(k) is not None else None
    m = OrcMesh()
    m.ncells, m.nint, m.nbnd = keep["V"].size, keep["owner"].size, keep["boundary_owner"].size
    m.owner, m.neighbour, m.bown = (keep[k].ctypes.data for k in ("owner", "neighbour", "boundary_owner"))
    m.V = keep["V"].ctypes.data
    m.Sf, m.C, m.weights = (keep[k].ctypes.data if keep[k] is not None else None for k in ("Sf", "C", "weights"))
    return m, keep


def transport_mules(mesh: dict, scheme: str, psi, phi, dt, psi_max, *, S=None, psi_b=None, psi_min=0.0, n_limiter_iter=3):
    """The call sequence of MOC::solveWithMULES (MOC.C:198-272) on psi [K][ncells]: high-order flux of
    `scheme`, MULES::limit per field against psi_max[k] / psi_min, limitSum over the fields, upwind
    flux + MULES::explicitSolve.  Returns the new fields."""
    m, keep = _orc_mesh(mesh)
    out = np.array(psi, dtype=np.float64, order="C", copy=True)
    K, nc = out.shape
    S = np.zeros_like(out) if S is None else np.ascontiguousarray(S, dtype=np.float64)
    pb = None if psi_b is None else np.ascontiguousarray(psi_b, dtype=np.float64)
    pmax = np.ascontiguousarray(np.broadcast_to(np.asarray(psi_max, dtype=np.float64), (K,)))
    fn = lib().orc_transport_mules
    fn.restype = C.c_int
    fn.argtypes = [C.POINTER(OrcMesh), C.c_int, C.c_int, C.c_int64, _dp, _dp, _dp, C.c_int64, _dp, _dp, C.c_double,
                   C.c_int, C.c_double]
    rc = fn(C.byref(m), DIV_SCHEMES[scheme], K, nc, _p(out), _p(np.ascontiguousarray(phi, dtype=np.float64)), _p(pb),
            0 if pb is None else pb.shape[1], _p(S), _p(pmax), float(psi_min), int(n_limiter_iter), float(dt))
    if rc != 0:
        raise RuntimeError(f"orc_transport_mules failed: {rc}")
    del keep
    return out


def div_flux(mesh: dict, scheme: str, phi, psi, psi_b=None):
    """fvc::flux(phi, psi, scheme) on every face [internal + boundary]"""
    m, keep = _orc_mesh(mesh)
    psi = np.ascontiguousarray(psi, dtype=np.float64)
    out = np.empty(m.nint + m.nbnd)
    fn = lib().orc_div_flux
    fn.restype = C.c_int
    fn.argtypes = [C.POINTER(OrcMesh), C.c_int, _dp, _dp, _dp, _dp]
    rc = fn(C.byref(m), DIV_SCHEMES[scheme], _p(np.ascontiguousarray(phi, dtype=np.float64)), _p(psi),
            _p(None if psi_b is None else np.ascontiguousarray(psi_b, dtype=np.float64)), _p(out))
    if rc != 0:
        raise RuntimeError(f"orc_div_flux failed: {rc}")
    del keep
    return out
