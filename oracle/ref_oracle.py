"""ctypes access to oracle/_ref/libpbe_ref.so -- TEST INFRASTRUCTURE.

libpbe_ref.so is the reference's OWN code: /root/reference/src/PBESystem/{PBEMethod,MOC,QMOM,Utility,
MomentInversion}.C and the kernel / daughter-distribution sources compiled UNMODIFIED against the
OpenFOAM / Eigen API shim of oracle/ref_shim (recipe: `make -C oracle ref`; built in the container
that has /root/reference, travels to the GPU box as a built file).  It pins oracle/pbe_oracle.c on
the reference itself (tests/test_ref_oracle.py) and is bench.py's `--impl reference` arm.

Only tests/, __graft_entry__ and bench.py's CPU legs may import this module.

Because the reference caches function-local `static const volScalarField&` in its kernels
(coalescenceKernels/CoulaloglouTavlarides.C:52-53) it is one-mesh-per-process and not thread safe;
`nthreads > 1` therefore runs one worker PROCESS per thread on disjoint cell blocks -- how the
reference itself goes parallel (decomposePar + mpirun, tests/validation/.../decomposeParDict:18-20).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libpbe_ref.so")
_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def available() -> bool:
    return os.path.exists(LIB_PATH)


def build(reference: str = "/root/reference") -> bool:
    """`make -C oracle ref` when the reference tree is present (this container); False otherwise."""
    if not os.path.isdir(os.path.join(reference, "src", "PBESystem")):
        return available()
    subprocess.check_call(["make", "-C", _HERE, "-s", "ref", f"REFERENCE={reference}"])
    return available()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB_PATH} is missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(LIB_PATH)
        L.ref_last_error.restype = C.c_char_p
        L.ref_method_steps.restype = C.c_int
        L.ref_method_steps.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp, _dp,
                                       C.c_double, C.c_int, _dp, _dp, _dp]
        L.ref_set_mesh.restype = C.c_int
        L.ref_set_mesh.argtypes = [C.c_int64, C.c_int64, _ip, _ip, C.c_int64, _ip, _dp, _dp, _dp, C.c_char_p, C.c_int]
        for name, n in (("ref_ct_coalescence_S", 10), ("ref_constant_coalescence_S", 1), ("ref_no_coalescence_S", 0),
                        ("ref_ct_breakup_S", 7), ("ref_binary_breakup_S", 1), ("ref_no_breakup_S", 1),
                        ("ref_uniform_binary_beta", 2)):
            fn = getattr(L, name)
            fn.restype = C.c_double
            fn.argtypes = [C.c_double] * n
        L.ref_uniform_binary_moment.restype = C.c_double
        L.ref_uniform_binary_moment.argtypes = [C.c_double, C.c_int]
        L.ref_wheeler.restype = C.c_int
        L.ref_wheeler.argtypes = [C.c_int, _dp, _dp, _dp]
        L.ref_set_quiet.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _arr(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


# ---- the dictionary a case's phaseProperties holds (tests/integration/MOC/diffs/phaseProperties.*) ------
def _coeffs(kind: str, name: str, c) -> str:
    if name in ("CoulaloglouTavlaridesC", "CoulaloglouTavlarides"):
        c1, c2, gamma, sigma = (float(x) for x in c)
        body = f"c1 {c1!r}; c2 {c2!r}; gamma {gamma!r}; sigma {sigma!r};"
    elif name == "constantCoalescence":
        body = f"constant {float(c[0])!r};"
    else:
        body = ""
    return f"{kind} {name}; {name}Coeffs {{ {body} }}\n"


def pbe_dict(method: str, coalescence: str, coal_coeffs, breakup: str, br_coeffs, *, N: int = 0, xi1: float = 0.0,
             relax: float = 1.0, order: int = 0, usingMULES: bool = True) -> str:
    txt = f"method {method};\n"
    if method == "methodOfClasses":
        txt += (f"MOCCoeffs {{ numberOfClasses {int(N)}; xi1 {float(xi1)!r}; usingMULES {'true' if usingMULES else 'false'}; "
                f"sourceRelaxation {float(relax)!r}; }}\n")
    elif method == "quadratureMethodOfMoments":
        txt += f"QMOMCoeffs {{ quadratureOrder {int(order)}; usingMULES {'true' if usingMULES else 'false'}; }}\n"
    txt += _coeffs("coalescenceKernel", coalescence, coal_coeffs)
    txt += _coeffs("breakupKernel", breakup, br_coeffs)
    txt += "daughterParticleDistribution uniformBinaryBreakup; uniformBinaryBreakupCoeffs { }\n"
    return txt


def method_steps(dict_text: str, stem: str, fields, eps, rho, nu, dt, *, alpha=None, vcell=None, phi=None,
                 nsteps: int = 1, want_sources: bool = True, want_d: bool = True):
    """nsteps x { method->correct(); method->d(); } of the reference.  Returns
    (fields after the steps, sources of the last step or None, d or None, seconds inside the loop)."""
    f = np.array(fields, dtype=np.float64, order="C", copy=True)
    K, nc = f.shape
    S = np.empty_like(f) if want_sources else None
    d = np.empty(nc) if want_d else None
    secs = C.c_double(0.0)
    eps, rho, nu, alpha, vcell, phi = (_arr(a) for a in (eps, rho, nu, alpha, vcell, phi))
    rc = lib().ref_method_steps(dict_text.encode(), stem.encode(), K, nc, _p(f), _p(eps), _p(rho), _p(nu), _p(alpha),
                                _p(vcell), _p(phi), float(dt), int(nsteps), _p(S), _p(d), C.byref(secs))
    if rc != 0:
        raise RuntimeError(f"reference failed ({rc}): {lib().ref_last_error().decode(errors='replace')}")
    return f, S, d, secs.value


def set_mesh(ncells, owner, neighbour, boundary_owner, *, Sf=None, C_=None, weights=None, div_scheme="Gauss upwind",
             n_limiter_iter=3):
    """Face addressing (+ geometry for limited schemes) of the process's single mesh; all-empty = 0-D cells."""
    own, nei, bown = (np.ascontiguousarray(a, dtype=np.int32) for a in (owner, neighbour, boundary_owner))
    Sf, C_, weights = (_arr(a) for a in (Sf, C_, weights))
    rc = lib().ref_set_mesh(int(ncells), own.size, own.ctypes.data_as(_ip), nei.ctypes.data_as(_ip), bown.size,
                            bown.ctypes.data_as(_ip), _p(Sf), _p(C_), _p(weights), div_scheme.encode(), int(n_limiter_iter))
    if rc != 0:
        raise RuntimeError("ref_set_mesh failed")


def clear_mesh():
    e = np.empty(0, dtype=np.int32)
    set_mesh(0, e, e, e)


# ---- one correct() of MOC::classSourceTerm for every class, in worker processes when nthreads > 1 ----------
def moc_source(coalescence, coal_coeffs, breakup, br_coeffs, N, xi1, relax, n, eps, rho, nu, dt, nthreads: int = 1,
               return_seconds: bool = False):
    """S [N][ncells] = the reference's classSourceTerm(k) for every k (MOC.C:87-157), from one
    MOC::correct() with zero flux.  nthreads > 1: one process per thread on disjoint cell blocks."""
    txt = pbe_dict("methodOfClasses", coalescence, coal_coeffs, breakup, br_coeffs, N=N, xi1=xi1, relax=relax)
    n = np.ascontiguousarray(n, dtype=np.float64)
    if nthreads <= 1:
        _, S, _, secs = method_steps(txt, "n", n, eps, rho, nu, dt, want_d=False)
        return (S, secs) if return_seconds else S
    S, secs = _parallel("n", txt, n, eps, rho, nu, dt, nthreads)
    return (S, secs) if return_seconds else S


def qmom_source(coalescence, coal_coeffs, breakup, br_coeffs, order, m, eps, rho, nu, dt=1.0, nthreads: int = 1,
                return_seconds: bool = False):
    """S [2 order][ncells] = the reference's momentSourceTerm(k) for every k after its NaN scrub
    (QMOM.C:88-110, 215-267; one wheeler_inversion per moment index), from one QMOM::correct()."""
    txt = pbe_dict("quadratureMethodOfMoments", coalescence, coal_coeffs, breakup, br_coeffs, order=order)
    m = np.ascontiguousarray(m, dtype=np.float64)
    if nthreads <= 1:
        _, S, _, secs = method_steps(txt, "m", m, eps, rho, nu, dt, want_d=False)
        return (S, secs) if return_seconds else S
    S, secs = _parallel("m", txt, m, eps, rho, nu, dt, nthreads)
    return (S, secs) if return_seconds else S


def _parallel(stem, txt, fields, eps, rho, nu, dt, nthreads, nsteps=1, warmup=0):
    """Split the cells over `nthreads` worker processes (python -m oracle.ref_oracle worker ...).  Every
    worker runs `warmup` untimed and then `nsteps` timed correct() calls (a time loop, as
    applications/PBEFoam/PBEFoam.C:56-67), all starting at a common wall-clock instant.  Returns the
    sources of the last step and the seconds per step of the slowest worker."""
    K, nc = fields.shape
    nthreads = max(1, min(int(nthreads), nc))
    bounds = np.linspace(0, nc, nthreads + 1).astype(np.int64)
    with tempfile.TemporaryDirectory(prefix="pbe_ref_") as tmp:
        start_at = time.time() + 2.0 + 0.02 * nthreads
        procs = []
        for r in range(nthreads):
            a, b = int(bounds[r]), int(bounds[r + 1])
            np.savez(os.path.join(tmp, f"in{r}.npz"), fields=fields[:, a:b], eps=np.asarray(eps)[a:b],
                     rho=np.asarray(rho)[a:b], nu=np.asarray(nu)[a:b])
            procs.append(subprocess.Popen([sys.executable, "-m", "oracle.ref_oracle", "worker", tmp, str(r), stem,
                                           repr(float(dt)), repr(start_at), str(int(nsteps)), str(int(warmup))],
                                          cwd=os.path.dirname(_HERE), stdin=subprocess.PIPE))
            procs[-1].stdin.write(txt.encode())
            procs[-1].stdin.close()
        for p in procs:
            if p.wait() != 0:
                raise RuntimeError("reference worker failed")
        S = np.empty_like(fields)
        secs = 0.0
        for r in range(nthreads):
            a, b = int(bounds[r]), int(bounds[r + 1])
            out = np.load(os.path.join(tmp, f"out{r}.npz"))
            S[:, a:b] = out["S"]
            secs = max(secs, float(out["secs"]))
    return S, secs / max(1, int(nsteps))


def _worker(argv):
    tmp, r, stem, dt, start_at = argv[0], int(argv[1]), argv[2], float(argv[3]), float(argv[4])
    nsteps, warmup = int(argv[5]), int(argv[6])
    txt = sys.stdin.read()
    d = np.load(os.path.join(tmp, f"in{r}.npz"))
    lib()
    delay = start_at - time.time()
    if delay > 0:
        time.sleep(delay)
    fields = d["fields"]
    if warmup > 0:
        fields, _, _, _ = method_steps(txt, stem, fields, d["eps"], d["rho"], d["nu"], dt, nsteps=warmup,
                                       want_sources=False, want_d=False)
    _, S, _, secs = method_steps(txt, stem, fields, d["eps"], d["rho"], d["nu"], dt, nsteps=nsteps, want_d=False)
    np.savez(os.path.join(tmp, f"out{r}.npz"), S=S, secs=secs)


def correct_timed(method, coalescence, coal_coeffs, breakup, br_coeffs, fields, eps, rho, nu, dt, *, N=0, xi1=0.0,
                  relax=1.0, order=0, nthreads=1, nsteps=1, warmup=0):
    """Seconds per correct() + d() of the reference (time loop of `nsteps` steps after `warmup` untimed
    ones) on these cells, one worker process per thread.  method: "methodOfClasses" |
    "quadratureMethodOfMoments"."""
    txt = pbe_dict(method, coalescence, coal_coeffs, breakup, br_coeffs, N=N, xi1=xi1, relax=relax, order=order)
    stem = "n" if method == "methodOfClasses" else "m"
    fields = np.ascontiguousarray(fields, dtype=np.float64)
    if nthreads <= 1:
        if warmup > 0:
            fields, _, _, _ = method_steps(txt, stem, fields, eps, rho, nu, dt, nsteps=warmup, want_sources=False,
                                           want_d=False)
        _, _, _, secs = method_steps(txt, stem, fields, eps, rho, nu, dt, nsteps=nsteps, want_sources=False, want_d=False)
        return secs / max(1, nsteps)
    _, secs = _parallel(stem, txt, fields, eps, rho, nu, dt, nthreads, nsteps=nsteps, warmup=warmup)
    return secs


# ---- scalar functors and the inversion ------------------------------------------------------------------------
def ct_coalescence_S(c, xi1, xi2, eps, rho, nu, V=1.0):
    return lib().ref_ct_coalescence_S(*[float(x) for x in c], float(xi1), float(xi2), float(eps), float(rho), float(nu),
                                      float(V))


def ct_breakup_S(c, xi, rho_d, eps):
    return lib().ref_ct_breakup_S(*[float(x) for x in c], float(xi), float(rho_d), float(eps))


def wheeler(m):
    m = np.ascontiguousarray(m, dtype=np.float64)
    x, w = np.empty(m.size // 2), np.empty(m.size // 2)
    lib().ref_wheeler(m.size, _p(m), _p(x), _p(w))
    return x, w


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "worker":
        _worker(sys.argv[2:])
    else:
        print("built" if build() else "not built")
