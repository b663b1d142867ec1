"""Thin Python handles over the C ABI, named after the reference's classes.

`MOC` / `QMOM` correspond to PBEMethods::MOC / PBEMethods::QMOM (src/PBESystem/MOC.H:57-127,
QMOM.H:55-112): constructed from the same coefficients (`numberOfClasses`, `xi1`,
`sourceRelaxation`, `quadratureOrder`) and kernel TypeNames; `source()` / `step()` are the
per-time-step work of `correct()` without the mesh-coupled transport.  Fields are torch CUDA
tensors, float64, shape [N, ncells] (class-major, as the reference's PtrList<volScalarField>).
All arithmetic happens inside libpbe_b200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import capi

__all__ = ["Context", "MOC", "QMOM", "MOM", "MOCState", "QMOMState", "Mesh", "make_spec", "kernel_types"]

_dp = C.POINTER(C.c_double)


def kernel_types(coalescence: str, breakup: str, daughter: str = "uniformBinaryBreakup"):
    """TypeName -> id through the library's run-time selection tables (raises on unknown)."""
    L = capi.lib()
    out = []
    for fn, name in ((L.pbe_coalescence_type, coalescence), (L.pbe_breakup_type, breakup),
                     (L.pbe_daughter_type, daughter)):
        v = C.c_int(-1)
        capi.check(fn(name.encode(), C.byref(v)))
        out.append(v.value)
    return tuple(out)


def make_spec(coalescence: str = "noCoalescence", coal_coeffs: Sequence[float] = (),
              breakup: str = "noBreakup", br_coeffs: Sequence[float] = (),
              daughter: str = "uniformBinaryBreakup",
              coal_table: Optional[np.ndarray] = None,
              br_table: Optional[np.ndarray] = None) -> capi.KernelSpec:
    """Kernel selection by the reference's TypeNames; `table` selects the user-kernel upload."""
    s = capi.KernelSpec()
    c_id = capi.COAL_TABLE if coalescence == "table" else None
    b_id = capi.BR_TABLE if breakup == "table" else None
    ids = kernel_types("noCoalescence" if c_id is not None else coalescence,
                       "noBreakup" if b_id is not None else breakup, daughter)
    s.coalescence = c_id if c_id is not None else ids[0]
    s.breakup = b_id if b_id is not None else ids[1]
    s.daughter = ids[2]
    for k, v in enumerate(coal_coeffs):
        s.coal_c[k] = float(v)
    for k, v in enumerate(br_coeffs):
        s.br_c[k] = float(v)
    keep = []
    if coal_table is not None:
        ct = np.ascontiguousarray(coal_table, dtype=np.float64)
        keep.append(ct)
        s.coal_table = ct.ctypes.data_as(_dp)
    if br_table is not None:
        bt = np.ascontiguousarray(br_table, dtype=np.float64)
        keep.append(bt)
        s.br_table = bt.ctypes.data_as(_dp)
    s._keep = keep
    return s


def _ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    assert t.dtype == torch.float64 and t.is_contiguous(), "fields are contiguous float64"
    return C.c_void_p(t.data_ptr())


def _ptr_rows(t: Optional[torch.Tensor]):
    """[rows][cols] float64 with unit column stride (rows may be padded): pointer to element (0, 0)."""
    if t is None:
        return None
    assert t.dtype == torch.float64 and t.dim() == 2 and (t.shape[1] <= 1 or t.stride(1) == 1)
    return C.c_void_p(t.data_ptr())


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Context:
    """pbe_ctx: one per (process, device)."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        self.device = int(device)
        self._deps = []  # weak references to the handles created on this context, in creation order
        capi.check(capi.lib().pbe_ctx_create(self.device, C.byref(self._h)))

    def _adopt(self, obj):
        """Handles (methods, meshes, resident states) hold a pointer to the C context: the C ABI wants them destroyed
        BEFORE it (include/pbe_b200.h).  close() therefore closes what is still alive first, newest first -- a handle
        that the garbage collector reaches after its context would otherwise read freed memory in its destructor."""
        import weakref
        self._deps.append(weakref.ref(obj))

    def close(self):
        if self._h:
            for r in reversed(self._deps):
                o = r()
                if o is not None:
                    o.close()
            self._deps = []
            capi.lib().pbe_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def info(self):
        name = C.create_string_buffer(128)
        sm = C.c_int()
        hbm = C.c_int64()
        capi.check(capi.lib().pbe_ctx_info(self._h, name, 128, C.byref(sm), C.byref(hbm)))
        return {"name": name.value.decode(), "sm_count": sm.value, "hbm_bytes": hbm.value}

    def launch_count(self) -> int:
        return int(capi.lib().pbe_ctx_launch_count(self._h))

    def fp64_peak_tflops(self):
        tf, ms = C.c_double(), C.c_double()
        capi.check(capi.lib().pbe_measure_fp64_peak(self._h, C.byref(tf), C.byref(ms)))
        return tf.value, ms.value

    # ---- scalar functors on the device (host numpy in / out) --------------------------------
    def eval_coalescence(self, spec, xi1, xi2, eps, rho, nu):
        a = [np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64) for v in (xi1, xi2, eps, rho, nu)]
        a = np.broadcast_arrays(*a)
        a = [np.ascontiguousarray(v) for v in a]
        out = np.empty_like(a[0])
        capi.check(capi.lib().pbe_eval_coalescence(
            self._h, C.byref(spec), out.size, *[v.ctypes.data_as(C.c_void_p) for v in a],
            out.ctypes.data_as(C.c_void_p)))
        return out

    def eval_breakup(self, spec, xi, rho, eps):
        a = [np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64) for v in (xi, rho, eps)]
        a = [np.ascontiguousarray(v) for v in np.broadcast_arrays(*a)]
        out = np.empty_like(a[0])
        capi.check(capi.lib().pbe_eval_breakup(
            self._h, C.byref(spec), out.size, *[v.ctypes.data_as(C.c_void_p) for v in a],
            out.ctypes.data_as(C.c_void_p)))
        return out

    def eval_daughter_beta(self, xi1, xi2, daughter=capi.DAUGHTER_UNIFORM_BINARY):
        a = [np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64) for v in (xi1, xi2)]
        a = [np.ascontiguousarray(v) for v in np.broadcast_arrays(*a)]
        out = np.empty_like(a[0])
        capi.check(capi.lib().pbe_eval_daughter_beta(
            self._h, daughter, out.size, a[0].ctypes.data_as(C.c_void_p),
            a[1].ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def eval_daughter_moment(self, xi, k, daughter=capi.DAUGHTER_UNIFORM_BINARY):
        xi = np.ascontiguousarray(np.atleast_1d(xi), dtype=np.float64)
        k = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(k), xi.shape), dtype=np.int32)
        out = np.empty_like(xi)
        capi.check(capi.lib().pbe_eval_daughter_moment(
            self._h, daughter, out.size, xi.ctypes.data_as(C.c_void_p),
            k.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def eval_integrate_exp(self, rate, lower):
        a = [np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64) for v in (rate, lower)]
        a = [np.ascontiguousarray(v) for v in np.broadcast_arrays(*a)]
        out = np.empty_like(a[0])
        capi.check(capi.lib().pbe_eval_integrate_exp(
            self._h, out.size, a[0].ctypes.data_as(C.c_void_p), a[1].ctypes.data_as(C.c_void_p),
            out.ctypes.data_as(C.c_void_p)))
        return out

    def wheeler_batch(self, m: torch.Tensor):
        """m [2*order, ncells] -> (x, w) each [order, ncells]"""
        nm, ncells = m.shape
        order = nm // 2
        x = torch.empty((order, ncells), dtype=torch.float64, device=m.device)
        w = torch.empty_like(x)
        capi.check(capi.lib().pbe_wheeler_batch(self._h, order, ncells, m.stride(0) if ncells else 0,
                                                _ptr(m), _ptr(x), _ptr(w), _stream_ptr(m.device)))
        return x, w

    def scrub_nan(self, f: torch.Tensor):
        """NaN -> 0 in place (QMOM.C:101-107)."""
        capi.check(capi.lib().pbe_field_scrub_nan(self.handle, f.numel(), _ptr(f), _stream_ptr(f.device)))
        return f

    def qmom_clamp_d(self, m: torch.Tensor, want_d: bool = True):
        """m_k = max(m_k, SMALL) in place and d = (6/pi m1/m0)^(1/3) (QMOM.C:117-125)."""
        K, ncells = m.shape
        d = torch.empty(ncells, dtype=torch.float64, device=m.device) if want_d else None
        capi.check(capi.lib().pbe_qmom_clamp_d(self.handle, K, ncells, m.stride(0) if ncells else 0, _ptr(m),
                                               _ptr(d), _stream_ptr(m.device)))
        return d

    def field_avg_max_min(self, f: torch.Tensor, vcell: Optional[torch.Tensor] = None):
        out = torch.empty(4, dtype=torch.float64, device=f.device)
        capi.check(capi.lib().pbe_field_avg_max_min(self._h, f.numel(), _ptr(f), _ptr(vcell),
                                                    _ptr(out), _stream_ptr(f.device)))
        return out


class MOC:
    """pbe_moc: PBEMethods::MOC's source path (MOCCoeffs: numberOfClasses, xi1, sourceRelaxation)."""

    def __init__(self, ctx: Context, numberOfClasses: int, xi1: float, spec: capi.KernelSpec,
                 sourceRelaxation: float = 1.0, pivots=None):
        """pivots: class volumes of a geometric / arbitrary grid (fixed-pivot extension, row f4; xi1 is
        then ignored); None: the reference's uniform grid (i + 1) xi1."""
        self.ctx = ctx
        ctx._adopt(self)
        self.N = int(numberOfClasses)
        self._h = C.c_void_p()
        if pivots is not None:
            pv = np.ascontiguousarray(pivots, dtype=np.float64)
            assert pv.shape == (self.N,)
            self.xi1 = float(pv[0])
            capi.check(capi.lib().pbe_moc_create_pivots(ctx.handle, self.N, pv.ctypes.data_as(C.POINTER(C.c_double)),
                                                        C.byref(spec), float(sourceRelaxation), C.byref(self._h)))
            return
        self.xi1 = float(xi1)
        capi.check(capi.lib().pbe_moc_create(ctx.handle, self.N, self.xi1, C.byref(spec),
                                             float(sourceRelaxation), C.byref(self._h)))

    def grid(self):
        """(class volumes [N], is_uniform)"""
        v = np.empty(self.N)
        u = C.c_int()
        capi.check(capi.lib().pbe_moc_grid(self._h, v.ctypes.data_as(C.POINTER(C.c_double)), C.byref(u)))
        return v, bool(u.value)

    def close(self):
        if self._h:
            capi.lib().pbe_moc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_kernel(self, which: str):
        """Pin the device kernel: "smem" or "tmem" (raises PbeError if unavailable for this N)."""
        ids = {"smem": capi.MOC_KERNEL_SMEM, "tmem": capi.MOC_KERNEL_TMEM}
        capi.check(capi.lib().pbe_moc_set_kernel(self._h, ids[which]))

    def kernel(self):
        w, b, c = C.c_int(), C.c_int(), C.c_int()
        capi.check(capi.lib().pbe_moc_get_kernel(self._h, C.byref(w), C.byref(b), C.byref(c)))
        return {"kernel": {capi.MOC_KERNEL_SMEM: "smem", capi.MOC_KERNEL_TMEM: "tmem",
                           capi.MOC_KERNEL_PIVOT: "pivot"}[w.value],
                "cells_per_cta": b.value, "ctas_per_sm": c.value}

    def _dims(self, n):
        assert n.dim() == 2 and n.shape[0] == self.N and n.stride(1) in (0, 1)
        ncells = n.shape[1]
        return ncells, (n.stride(0) if ncells else 0)

    def source(self, n, eps, rho, nu, dt, out=None):
        ncells, stride = self._dims(n)
        S = torch.empty_like(n) if out is None else out
        capi.check(capi.lib().pbe_moc_source(self._h, ncells, stride, _ptr(n), _ptr(eps), _ptr(rho),
                                             _ptr(nu), float(dt), _ptr(S), _stream_ptr(n.device)))
        return S

    def step(self, n, eps, rho, nu, dt, *, S=None, d=None, alpha=None, vcell=None, moments=None,
             cont_err_max=None):
        """In-place n += dt S with the requested fused outputs."""
        ncells, stride = self._dims(n)
        io = capi.MocIO()
        io.alpha = _ptr(alpha)
        io.vcell = _ptr(vcell)
        io.S = _ptr(S)
        io.d = _ptr(d)
        io.moments = _ptr(moments)
        io.n_moments = 0 if moments is None else moments.numel()
        io.cont_err_max = _ptr(cont_err_max)
        capi.check(capi.lib().pbe_moc_step(self._h, ncells, stride, _ptr(n), _ptr(eps), _ptr(rho),
                                           _ptr(nu), float(dt), C.byref(io), _stream_ptr(n.device)))
        return n

    def d(self, n):
        ncells, stride = self._dims(n)
        d = torch.empty(ncells, dtype=torch.float64, device=n.device)
        capi.check(capi.lib().pbe_moc_d(self._h, ncells, stride, _ptr(n), _ptr(d), _stream_ptr(n.device)))
        return d

    def class_minmax(self, n):
        ncells, stride = self._dims(n)
        mn = torch.empty(self.N, dtype=torch.float64, device=n.device)
        mx = torch.empty_like(mn)
        capi.check(capi.lib().pbe_moc_class_minmax(self._h, ncells, stride, _ptr(n), _ptr(mn), _ptr(mx),
                                                   _stream_ptr(n.device)))
        return mn, mx

    def psi_max(self, alpha, vcell):
        """max_c(alpha V / xi_k) per class (MOC.C:238): the upper bound MULES::limit is given"""
        out = torch.empty(self.N, dtype=torch.float64, device=alpha.device)
        capi.check(capi.lib().pbe_moc_psi_max(self._h, alpha.numel(), _ptr(alpha), _ptr(vcell), _ptr(out),
                                              _stream_ptr(alpha.device)))
        return out

    def source_host(self, n, eps, rho, nu, dt, out=None):
        """Host-buffer entry point (numpy or CPU torch tensors; pinned memory gives full rate)."""
        n_, eps_, rho_, nu_ = (_as_host(v) for v in (n, eps, rho, nu))
        assert n_.shape[0] == self.N
        S = torch.empty_like(n_) if out is None else _as_host(out)
        capi.check(capi.lib().pbe_moc_source_host(
            self._h, n_.shape[1], n_.stride(0) if n_.shape[1] else 0, _ptr(n_), _ptr(eps_), _ptr(rho_),
            _ptr(nu_), float(dt), _ptr(S)))
        return S


class _CudaBuf:
    """A raw device pointer with the __cuda_array_interface__ torch.as_tensor understands: zero-copy
    views of memory the library owns (the resident states)."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": "<f8",
                                         "data": (int(ptr), False), "version": 2, "strides": None}


def _device_view(ptr: int, shape, device: int):
    if any(int(x) == 0 for x in shape):
        return torch.empty(tuple(shape), dtype=torch.float64, device=f"cuda:{device}")
    return torch.as_tensor(_CudaBuf(ptr, shape), device=f"cuda:{device}")


def _hptr(t):
    """host tensor (or None) -> void* for the resident entries"""
    return None if t is None else C.c_void_p(_as_host(t).data_ptr())


def _fvm_controls(tolerance=1e-5, rel_tol=0.0, min_iter=1, max_iter=1000, relaxation=0.0, solver="auto"):
    return capi.FvmControls(float(tolerance), float(rel_tol), int(min_iter), int(max_iter), float(relaxation),
                            {"auto": 0, "bicgstab": 1, "jacobi": 2}[solver] if isinstance(solver, str) else int(solver))


class MOCState:
    """pbe_moc_state: the class fields owned on the device, as PBEMethods::MOC owns
    classNumberDensity_ (MOC.H:66, MOC.C:69-77); `correct()` is one MOC::correct() + d() with only
    this step's phase fields crossing the bus (PBEDiameter.H:95-99)."""

    def __init__(self, moc: "MOC", ncells: int, vcell=None, n0=None, mesh: "Mesh" = None, scheme="upwind"):
        self.moc, self.ncells = moc, int(ncells)
        moc.ctx._adopt(self)  # created after its method and mesh: closed before them
        self._h = C.c_void_p()
        self._mesh = mesh  # keep alive
        capi.check(capi.lib().pbe_moc_state_create(moc._h, self.ncells, _hptr(vcell), C.byref(self._h)))
        if n0 is not None:
            self.upload(n0)
        if mesh is not None:
            capi.check(capi.lib().pbe_moc_state_attach_mesh(self._h, mesh._h, div_scheme(scheme)))

    def set_implicit(self, on=True, **controls):
        """usingMULES false (MOC.C:58, 177-180): correct() solves fvm::ddt + fvm::div == S (solveWithFVM) instead of
        the MULES update; controls as fvSolution's (tolerance, rel_tol, min_iter, max_iter, relaxation, solver)."""
        ctl = _fvm_controls(**controls) if on else None
        capi.check(capi.lib().pbe_moc_state_set_implicit(self._h, C.byref(ctl) if on else None))

    def solver_performance(self):
        """[(initial residual, final residual, iterations, state)] per class of the last implicit step"""
        perf = (capi.FvmPerf * self.moc.N)()
        capi.check(capi.lib().pbe_moc_state_solver_performance(self._h, C.cast(perf, C.c_void_p)))
        return [(p.initial_residual, p.final_residual, p.iterations, p.state) for p in perf]

    def close(self):
        if self._h:
            capi.lib().pbe_moc_state_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, n):
        n_ = _as_host(n)
        assert n_.shape == (self.moc.N, self.ncells)
        capi.check(capi.lib().pbe_moc_state_upload(self._h, n_.stride(0) if self.ncells else 0, _ptr_rows(n_)))

    def download(self, out=None):
        out = torch.empty((self.moc.N, self.ncells), dtype=torch.float64) if out is None else _as_host(out)
        capi.check(capi.lib().pbe_moc_state_download(self._h, out.stride(0) if self.ncells else 0, _ptr_rows(out)))
        return out

    def upload_range(self, cell0: int, n):
        """n [N][count]: the class fields of cells [cell0, cell0 + count)"""
        n_ = _as_host(n)
        capi.check(capi.lib().pbe_moc_state_upload_range(self._h, int(cell0), n_.shape[1],
                                                         n_.stride(0) if n_.shape[1] else 0, _ptr_rows(n_)))

    def download_range(self, cell0: int, count: int):
        out = torch.empty((self.moc.N, int(count)), dtype=torch.float64)
        capi.check(capi.lib().pbe_moc_state_download_range(self._h, int(cell0), int(count),
                                                           out.stride(0) if count else 0, _ptr_rows(out)))
        return out

    def device_fields(self):
        """zero-copy torch views of the state's device memory: (n [N][ncells], d [ncells])"""
        n, stride, d = C.c_void_p(), C.c_int64(), C.c_void_p()
        capi.check(capi.lib().pbe_moc_state_device(self._h, C.byref(n), C.byref(stride), C.byref(d)))
        assert stride.value == self.ncells
        dev = self.moc.ctx.device
        return (_device_view(n.value, (self.moc.N, self.ncells), dev), _device_view(d.value, (self.ncells,), dev))

    def d(self):
        d = torch.empty(self.ncells, dtype=torch.float64)
        capi.check(capi.lib().pbe_moc_state_d(self._h, _ptr(d)))
        return d

    def correct(self, eps, rho, nu, alpha, dt, *, phi=None, psi_b=None, d_out=None, want_d=True,
                minmax=False, n_moments=0):
        """One MOC::correct() + d(): host eps, rho, nu, alpha in; returns (d, diag dict)."""
        d = d_out if d_out is not None else (torch.empty(self.ncells, dtype=torch.float64) if want_d else None)
        diag = capi.MocDiag()
        mn = mx = None
        if minmax:
            mn, mx = np.empty(self.moc.N), np.empty(self.moc.N)
            diag.class_min, diag.class_max = mn.ctypes.data, mx.ctypes.data
        diag.n_moments = int(n_moments)
        pb = None if psi_b is None else _as_host(psi_b)
        capi.check(capi.lib().pbe_moc_correct_resident(
            self._h, _hptr(eps), _hptr(rho), _hptr(nu), _hptr(alpha), _hptr(phi), _hptr(pb),
            0 if pb is None else pb.stride(0), float(dt), _hptr(d), C.byref(diag)))
        return d, {"class_min": mn, "class_max": mx, "cont_err_max": diag.cont_err_max,
                   "moments": list(diag.moments)[:int(n_moments)]}


class QMOMState:
    """pbe_qmom_state: the moment fields owned on the device (QMOM.H:62, QMOM.C:72-79);
    `correct()` is one QMOM::correct() (QMOM.C:85-129)."""

    def __init__(self, q: "QMOM", ncells: int, vcell=None, m0=None, mesh: "Mesh" = None, scheme="upwind"):
        self.q, self.ncells = q, int(ncells)
        q.ctx._adopt(self)
        self._h = C.c_void_p()
        self._mesh = mesh
        capi.check(capi.lib().pbe_qmom_state_create(q._h, self.ncells, _hptr(vcell), C.byref(self._h)))
        if m0 is not None:
            self.upload(m0)
        if mesh is not None:
            capi.check(capi.lib().pbe_qmom_state_attach_mesh(self._h, mesh._h, div_scheme(scheme)))

    def set_implicit(self, on=True, **controls):
        """usingMULES false (QMOM.C:71, 112-115): correct() solves fvm::ddt + fvm::div == S (solveWithFVM) instead of
        the MULES update; controls as fvSolution's (tolerance, rel_tol, min_iter, max_iter, relaxation, solver)."""
        ctl = _fvm_controls(**controls) if on else None
        capi.check(capi.lib().pbe_qmom_state_set_implicit(self._h, C.byref(ctl) if on else None))

    def solver_performance(self):
        """[(initial residual, final residual, iterations, state)] per moment of the last implicit step"""
        perf = (capi.FvmPerf * (2 * self.q.order))()
        capi.check(capi.lib().pbe_qmom_state_solver_performance(self._h, C.cast(perf, C.c_void_p)))
        return [(p.initial_residual, p.final_residual, p.iterations, p.state) for p in perf]

    def close(self):
        if self._h:
            capi.lib().pbe_qmom_state_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, m):
        m_ = _as_host(m)
        assert m_.shape == (2 * self.q.order, self.ncells)
        capi.check(capi.lib().pbe_qmom_state_upload(self._h, m_.stride(0) if self.ncells else 0, _ptr_rows(m_)))

    def download(self, out=None):
        out = torch.empty((2 * self.q.order, self.ncells), dtype=torch.float64) if out is None else _as_host(out)
        capi.check(capi.lib().pbe_qmom_state_download(self._h, out.stride(0) if self.ncells else 0, _ptr_rows(out)))
        return out

    def d(self):
        d = torch.empty(self.ncells, dtype=torch.float64)
        capi.check(capi.lib().pbe_qmom_state_d(self._h, _ptr(d)))
        return d

    def device_fields(self):
        """zero-copy torch views of the state's device memory: (m [2 order][ncells], d [ncells])"""
        m, stride, d, S = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_void_p()
        capi.check(capi.lib().pbe_qmom_state_device(self._h, C.byref(m), C.byref(stride), C.byref(d), C.byref(S)))
        dev = self.q.ctx.device
        return (_device_view(m.value, (2 * self.q.order, self.ncells), dev), _device_view(d.value, (self.ncells,), dev))

    def correct(self, eps, rho, nu, dt, *, phi=None, psi_b=None, d_out=None, want_d=True, stats=False):
        """One QMOM::correct(); returns (d, stats dict of [rows][4] = sum f V, sum V, max, min)."""
        d = d_out if d_out is not None else (torch.empty(self.ncells, dtype=torch.float64) if want_d else None)
        diag, out = None, {}
        if stats:
            nm = 2 * self.q.order
            out = {"source": np.empty((nm, 4)), "moment": np.empty((nm, 4)), "d": np.empty(4)}
            diag = capi.QmomDiag()
            diag.source, diag.moment, diag.d = (out[k].ctypes.data for k in ("source", "moment", "d"))
        pb = None if psi_b is None else _as_host(psi_b)
        capi.check(capi.lib().pbe_qmom_correct_resident(
            self._h, _hptr(eps), _hptr(rho), _hptr(nu), _hptr(phi), _hptr(pb), 0 if pb is None else pb.stride(0),
            float(dt), _hptr(d), None if diag is None else C.byref(diag)))
        return d, out


def _as_host(v):
    if isinstance(v, np.ndarray):
        v = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64))
    assert v.device.type == "cpu" and v.dtype == torch.float64
    return v


class QMOM:
    """pbe_qmom: PBEMethods::QMOM's source path (QMOMCoeffs: quadratureOrder)."""

    def __init__(self, ctx: Context, quadratureOrder: int, spec: capi.KernelSpec):
        self.ctx = ctx
        ctx._adopt(self)
        self.order = int(quadratureOrder)
        self._h = C.c_void_p()
        capi.check(capi.lib().pbe_qmom_create(ctx.handle, self.order, C.byref(spec), C.byref(self._h)))

    def close(self):
        if self._h:
            capi.lib().pbe_qmom_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dims(self, m):
        assert m.dim() == 2 and m.shape[0] == 2 * self.order
        ncells = m.shape[1]
        return ncells, (m.stride(0) if ncells else 0)

    def source(self, m, eps, rho, nu, out=None):
        ncells, stride = self._dims(m)
        S = torch.empty_like(m) if out is None else out
        capi.check(capi.lib().pbe_qmom_source(self._h, ncells, stride, _ptr(m), _ptr(eps), _ptr(rho),
                                              _ptr(nu), _ptr(S), _stream_ptr(m.device)))
        return S

    def step(self, m, eps, rho, nu, dt, *, S=None, d=None):
        ncells, stride = self._dims(m)
        capi.check(capi.lib().pbe_qmom_step(self._h, ncells, stride, _ptr(m), _ptr(eps), _ptr(rho),
                                            _ptr(nu), float(dt), _ptr(S), _ptr(d),
                                            _stream_ptr(m.device)))
        return m

    def source_host(self, m, eps, rho, nu, out=None):
        m_, eps_, rho_, nu_ = (_as_host(v) for v in (m, eps, rho, nu))
        S = torch.empty_like(m_) if out is None else _as_host(out)
        capi.check(capi.lib().pbe_qmom_source_host(
            self._h, m_.shape[1], m_.stride(0) if m_.shape[1] else 0, _ptr(m_), _ptr(eps_), _ptr(rho_),
            _ptr(nu_), _ptr(S)))
        return S


class MOM:
    """pbe_mom: PBEMethods::MOM's per-cell path (methodOfMoments, MOM.C:148-236, 259-362): three
    moments, presumed gamma density, nested integrals on a fixed-depth Gauss-Kronrod rule."""

    def __init__(self, ctx: Context, spec: capi.KernelSpec, panels: int = 0):
        self.ctx = ctx
        ctx._adopt(self)
        self._h = C.c_void_p()
        capi.check(capi.lib().pbe_mom_create(ctx.handle, int(panels), C.byref(spec), C.byref(self._h)))

    def close(self):
        if self._h:
            capi.lib().pbe_mom_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def coverage(self, reset: bool = True):
        """(cells the fixed rule did not cover since the last reset, worst relative deviation of m0 / m1)"""
        n, w = C.c_int64(), C.c_double()
        capi.check(capi.lib().pbe_mom_coverage(self._h, C.byref(n), C.byref(w), int(reset)))
        return n.value, w.value

    @staticmethod
    def _dims(m):
        assert m.dim() == 2 and m.shape[0] == 3
        ncells = m.shape[1]
        return ncells, (m.stride(0) if ncells else 0)

    def source(self, m, eps, rho, nu, out=None, gamma=None):
        ncells, stride = self._dims(m)
        S = torch.empty_like(m) if out is None else out
        capi.check(capi.lib().pbe_mom_source(self._h, ncells, stride, _ptr(m), _ptr(eps), _ptr(rho),
                                             _ptr(nu), _ptr(S), _ptr(gamma), _stream_ptr(m.device)))
        return S

    def step(self, m, eps, rho, nu, dt, *, S=None, d=None, gamma=None):
        ncells, stride = self._dims(m)
        capi.check(capi.lib().pbe_mom_step(self._h, ncells, stride, _ptr(m), _ptr(eps), _ptr(rho),
                                           _ptr(nu), float(dt), _ptr(S), _ptr(d), _ptr(gamma),
                                           _stream_ptr(m.device)))
        return m


def div_scheme(name: str) -> int:
    """fvSchemes `div(phi,n.*)` entry -> PBE_DIV_*; unknown / unsupported names raise PbeError."""
    out = C.c_int()
    capi.check(capi.lib().pbe_div_scheme_type(name.encode(), C.byref(out)))
    return out.value


class Mesh:
    """pbe_mesh: OpenFOAM face addressing (owner / neighbour / boundary faceCells, V) on the device,
    and the explicit transport + source update of MOC::solveWithMULES (MOC.C:198-272) /
    QMOM::solveWithMULES (QMOM.C:131-197) for `Gauss upwind`."""

    def __init__(self, ctx: Context, owner, neighbour, boundary_owner, vcell):
        self.ctx = ctx
        ctx._adopt(self)
        own = np.ascontiguousarray(owner, dtype=np.int32)
        nei = np.ascontiguousarray(neighbour, dtype=np.int32)
        bown = np.ascontiguousarray(boundary_owner, dtype=np.int32)
        V = np.ascontiguousarray(vcell, dtype=np.float64)
        assert own.shape == nei.shape and own.ndim == 1 and bown.ndim == 1 and V.ndim == 1
        self.ncells, self.n_internal, self.n_boundary = V.size, own.size, bown.size
        self._h = C.c_void_p()
        capi.check(capi.lib().pbe_mesh_create(
            ctx.handle, V.size, own.size, own.ctypes.data, nei.ctypes.data, bown.size,
            bown.ctypes.data, V.ctypes.data, C.byref(self._h)))

    def close(self):
        if self._h:
            capi.lib().pbe_mesh_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int()
        capi.check(capi.lib().pbe_mesh_info(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        F, nreg = C.c_int(), C.c_int64()
        capi.check(capi.lib().pbe_mesh_regular_cells(self._h, C.byref(F), C.byref(nreg)))
        return {"ncells": a.value, "internal_faces": b.value, "boundary_faces": c.value,
                "max_faces_per_cell": d.value, "faces_per_regular_cell": F.value,
                "regular_cells": nreg.value}

    def moc_correct_host(self, moc: "MOC", n, eps, rho, nu, phi, dt, *, psi_b=None, want_d=True,
                         scheme="upwind"):
        """Host-buffer MOC::correct() with solveWithMULES: n [N][ncells] (CPU tensor / numpy) is
        updated in place; returns d (CPU tensor) or None."""
        n_, eps_, rho_, nu_, phi_ = (_as_host(v) for v in (n, eps, rho, nu, phi))
        pb = None if psi_b is None else _as_host(psi_b)
        d = torch.empty(n_.shape[1], dtype=torch.float64) if want_d else None
        capi.check(capi.lib().pbe_moc_correct_host(
            moc._h, self._h, div_scheme(scheme), n_.stride(0), _ptr(n_), _ptr(eps_), _ptr(rho_), _ptr(nu_),
            _ptr(phi_), _ptr(pb), 0 if pb is None else pb.stride(0), float(dt), _ptr(d)))
        return d

    def set_geometry(self, Sf, C_, weights):
        """fvMesh::Sf() [faces][3], C() [ncells][3], weights() [internal faces] for the limited schemes"""
        a = [np.ascontiguousarray(v, dtype=np.float64) for v in (Sf, C_, weights)]
        assert a[0].shape == (self.n_internal + self.n_boundary, 3) and a[1].shape == (self.ncells, 3)
        assert a[2].shape == (self.n_internal,)
        capi.check(capi.lib().pbe_mesh_set_geometry(self._h, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data))

    def transport_mules(self, psi, phi, dt, psi_max, *, scheme="vanLeer", psi_b=None, S=None, out=None, psi_min=0.0,
                        n_limiter_iter=3):
        """MOC::solveWithMULES's transport (MOC.C:198-272): limited high-order correction + upwind,
        limitSum over the fields, MULES::explicitSolve.  psi_max: device tensor [K]."""
        K, ncells = psi.shape
        out = torch.empty_like(psi) if out is None else out
        capi.check(capi.lib().pbe_transport_mules(
            self._h, div_scheme(scheme) if isinstance(scheme, str) else int(scheme), K, psi.stride(0) if ncells else 0,
            _ptr(psi), _ptr(phi), _ptr(psi_b), 0 if psi_b is None else psi_b.stride(0), _ptr(S), _ptr(psi_max),
            float(psi_min), int(n_limiter_iter), float(dt), _ptr(out), _stream_ptr(psi.device)))
        return out

    def transport_implicit(self, psi, phi, dt, *, scheme="upwind", psi_b=None, S=None, out=None, tolerance=1e-5,
                           rel_tol=0.0, min_iter=1, max_iter=1000, relaxation=0.0, solver="auto"):
        """MOC::solveWithFVM / QMOM::solveWithFVM (MOC.C:186-198, QMOM.C:201-209): fvm::ddt + fvm::div == S for all K
        fields at once.  Returns (psi_new, perf) with perf = [(initial residual, final residual, iterations, state)]."""
        K, ncells = psi.shape
        out = torch.empty_like(psi) if out is None else out
        ctl = capi.FvmControls(float(tolerance), float(rel_tol), int(min_iter), int(max_iter), float(relaxation),
                               {"auto": 0, "bicgstab": 1, "jacobi": 2}[solver] if isinstance(solver, str) else int(solver))
        perf = (capi.FvmPerf * K)()
        capi.check(capi.lib().pbe_transport_implicit(
            self._h, div_scheme(scheme) if isinstance(scheme, str) else int(scheme), K, psi.stride(0) if ncells else 0,
            _ptr(psi), _ptr(phi), _ptr(psi_b), 0 if psi_b is None else psi_b.stride(0), _ptr(S), float(dt),
            C.byref(ctl), _ptr(out), C.cast(perf, C.c_void_p), _stream_ptr(psi.device)))
        return out, [(p.initial_residual, p.final_residual, p.iterations, p.state) for p in perf]

    def set_halo_plan(self, neighbours, send_cells, global_ncells: int):
        """The decomposed implicit solve (pbe_mesh_set_halo_plan): neighbours = [(rank, first coupled face, count)] as
        sharding.partition_mesh gives them, send_cells = this rank's cell on every coupled face."""
        import numpy as np
        peer = np.ascontiguousarray([q[0] for q in neighbours], dtype=np.int32)
        off = np.ascontiguousarray([q[1] for q in neighbours], dtype=np.int64)
        cnt = np.ascontiguousarray([q[2] for q in neighbours], dtype=np.int64)
        sc = np.ascontiguousarray(send_cells, dtype=np.int32)
        capi.check(capi.lib().pbe_mesh_set_halo_plan(self._h, len(neighbours), peer.ctypes.data, off.ctypes.data,
                                                     cnt.ctypes.data, sc.ctypes.data, int(global_ncells)))

    def set_coupled(self, first_boundary_face: int, count: int):
        """Mark boundary faces [first, first+count) as processor-patch faces (decomposed mesh)."""
        capi.check(capi.lib().pbe_mesh_set_coupled(self._h, int(first_boundary_face), int(count)))
        self.n_coupled = int(count)

    def gather_cells(self, psi, cells, out=None):
        """out[k][j] = psi[k][cells[j]] (halo packing); cells: int32 device tensor."""
        K, ncells = psi.shape
        n = cells.numel()
        assert cells.dtype == torch.int32 and cells.is_contiguous() and cells.device == psi.device
        out = torch.empty((K, n), dtype=psi.dtype, device=psi.device) if out is None else out
        capi.check(capi.lib().pbe_gather_cells(self.ctx.handle, K, psi.stride(0) if ncells else 0, _ptr(psi), n,
                                               C.c_void_p(cells.data_ptr()), _ptr(out), out.stride(0) if n else 0,
                                               _stream_ptr(psi.device)))
        return out

    def transport_explicit(self, psi, phi, dt, *, psi_b=None, S=None, out=None, scheme="upwind",
                           psi_halo=None, cells=capi.CELLS_ALL):
        """psi [K][ncells] -> (psi rDeltaT + S - div(phi psi_upwind)) / rDeltaT, new tensor or `out`."""
        assert psi.dim() == 2
        K, ncells = psi.shape
        out = torch.empty_like(psi) if out is None else out
        capi.check(capi.lib().pbe_transport_explicit_part(
            self._h, div_scheme(scheme) if isinstance(scheme, str) else int(scheme), int(cells), K,
            psi.stride(0) if ncells else 0, _ptr(psi), _ptr(phi), _ptr(psi_b),
            0 if psi_b is None else psi_b.stride(0), _ptr_rows(psi_halo),
            0 if psi_halo is None else psi_halo.stride(0), _ptr(S), float(dt), _ptr(out),
            _stream_ptr(psi.device)))
        return out
