"""ctypes binding of libpbe_b200.so (C ABI: include/pbe_b200.h).

This is plumbing for tests/, bench.py and the Python tools: device memory and streams come from
PyTorch tensors (`tensor.data_ptr()`), everything that computes is a hand-written sm_100a kernel
inside the shared library.  There is no CPU fallback: if the library is missing or a call fails
the error is raised, never papered over.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libpbe_b200.so")

PBE_OK, PBE_E_INVALID, PBE_E_UNKNOWN_TYPE, PBE_E_CUDA, PBE_E_UNSUPPORTED, PBE_E_NOMEM = range(6)
COAL_NONE, COAL_CONSTANT, COAL_CT, COAL_TABLE = 0, 1, 2, 3
BR_NONE, BR_BINARY, BR_CT, BR_TABLE = 0, 1, 2, 3
DAUGHTER_UNIFORM_BINARY = 0
QMOM_MAX_ORDER = 4
MOC_KERNEL_SMEM, MOC_KERNEL_TMEM, MOC_KERNEL_PIVOT = 1, 2, 3
DIV_UPWIND, DIV_VANLEER, DIV_LIMITEDLINEAR = 0, 1, 2
CELLS_ALL, CELLS_REGULAR, CELLS_IRREGULAR = 0, 1, 2
FVM_SOLVER_AUTO, FVM_SOLVER_BICGSTAB, FVM_SOLVER_JACOBI = 0, 1, 2

_dp = C.POINTER(C.c_double)
_vp = C.c_void_p
_i64 = C.c_int64


class KernelSpec(C.Structure):
    """struct pbe_kernel_spec"""

    _fields_ = [
        ("coalescence", C.c_int),
        ("coal_c", C.c_double * 4),
        ("breakup", C.c_int),
        ("br_c", C.c_double * 4),
        ("daughter", C.c_int),
        ("coal_table", _dp),
        ("br_table", _dp),
    ]


class MocIO(C.Structure):
    """struct pbe_moc_io (device pointers)"""

    _fields_ = [
        ("alpha", _vp),
        ("vcell", _vp),
        ("S", _vp),
        ("d", _vp),
        ("moments", _vp),
        ("n_moments", C.c_int),
        ("cont_err_max", _vp),
    ]


class MocDiag(C.Structure):
    """struct pbe_moc_diag (host pointers)"""

    _fields_ = [
        ("class_min", _vp),
        ("class_max", _vp),
        ("cont_err_max", C.c_double),
        ("n_moments", C.c_int),
        ("moments", C.c_double * 8),
    ]


class QmomDiag(C.Structure):
    """struct pbe_qmom_diag (host pointers)"""

    _fields_ = [("source", _vp), ("moment", _vp), ("d", _vp)]


class PbeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libpbe_b200 error {code}: {msg}")
        self.code = code
        self.message = msg


# every exported symbol of include/pbe_b200.h: name -> (restype, argtypes)
class FvmControls(C.Structure):
    """pbe_fvm_controls (include/pbe_b200.h): fvSolution's tolerance / relTol / minIter / maxIter + the equation's
    relaxation factor."""
    _fields_ = [("tolerance", C.c_double), ("rel_tol", C.c_double), ("min_iter", C.c_int), ("max_iter", C.c_int),
                ("relaxation", C.c_double), ("solver", C.c_int)]


class FvmPerf(C.Structure):
    _fields_ = [("initial_residual", C.c_double), ("final_residual", C.c_double), ("iterations", C.c_int),
                ("state", C.c_int)]


PROTOTYPES = {
    "pbe_coalescence_type": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "pbe_breakup_type": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "pbe_daughter_type": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "pbe_last_error": (C.c_char_p, []),
    "pbe_version": (C.c_int, []),
    "pbe_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "pbe_ctx_destroy": (C.c_int, [_vp]),
    "pbe_ctx_info": (C.c_int, [_vp, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(_i64)]),
    "pbe_ctx_launch_count": (_i64, [_vp]),
    "pbe_measure_fp64_peak": (C.c_int, [_vp, _dp, _dp]),
    "pbe_moc_create": (C.c_int, [_vp, C.c_int, C.c_double, C.POINTER(KernelSpec), C.c_double,
                                 C.POINTER(_vp)]),
    "pbe_moc_create_pivots": (C.c_int, [_vp, C.c_int, _dp, C.POINTER(KernelSpec), C.c_double, C.POINTER(_vp)]),
    "pbe_moc_grid": (C.c_int, [_vp, _dp, C.POINTER(C.c_int)]),
    "pbe_moc_grid_device": (C.c_int, [_vp, C.POINTER(_vp)]),
    "pbe_moc_destroy": (C.c_int, [_vp]),
    "pbe_moc_set_kernel": (C.c_int, [_vp, C.c_int]),
    "pbe_moc_get_kernel": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pbe_moc_source": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp]),
    "pbe_moc_step": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.c_double, C.POINTER(MocIO),
                               _vp]),
    "pbe_moc_d": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "pbe_moc_class_minmax": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "pbe_moc_class_minmax_noalloc": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "pbe_moc_source_host": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.c_double, _vp]),
    "pbe_qmom_create": (C.c_int, [_vp, C.c_int, C.POINTER(KernelSpec), C.POINTER(_vp)]),
    "pbe_qmom_destroy": (C.c_int, [_vp]),
    "pbe_qmom_source": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "pbe_qmom_step": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp, _vp]),
    "pbe_wheeler_batch": (C.c_int, [_vp, C.c_int, _i64, _i64, _vp, _vp, _vp, _vp]),
    "pbe_qmom_source_host": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "pbe_mom_create": (C.c_int, [_vp, C.c_int, C.POINTER(KernelSpec), C.POINTER(_vp)]),
    "pbe_mom_destroy": (C.c_int, [_vp]),
    "pbe_mom_coverage": (C.c_int, [_vp, C.POINTER(_i64), _dp, C.c_int]),
    "pbe_mom_source": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "pbe_mom_step": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp, _vp, _vp]),
    "pbe_mesh_create": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp, C.POINTER(_vp)]),
    "pbe_mesh_destroy": (C.c_int, [_vp]),
    "pbe_mesh_info": (C.c_int, [_vp, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64),
                                C.POINTER(C.c_int)]),
    "pbe_mesh_regular_cells": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(_i64)]),
    "pbe_div_scheme_type": (C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    "pbe_transport_explicit": (C.c_int, [_vp, C.c_int, C.c_int, _i64, _vp, _vp, _vp, _i64, _vp,
                                         C.c_double, _vp, _vp]),
    "pbe_mesh_set_geometry": (C.c_int, [_vp, _vp, _vp, _vp]),
    "pbe_mesh_set_halo_plan": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _i64]),
    "pbe_transport_implicit": (C.c_int, [_vp, C.c_int, C.c_int, _i64, _vp, _vp, _vp, _i64, _vp, C.c_double, _vp, _vp,
                                          _vp, _vp]),
    "pbe_transport_mules": (C.c_int, [_vp, C.c_int, C.c_int, _i64, _vp, _vp, _vp, _i64, _vp, _vp, C.c_double,
                                      C.c_int, C.c_double, _vp, _vp]),
    "pbe_moc_psi_max": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "pbe_qmom_psi_max": (C.c_int, [_vp, C.c_int, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _vp]),
    "pbe_moc_num_classes": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(_vp)]),
    "pbe_moc_coeffs": (C.c_int, [_vp, _dp, _dp]),
    "pbe_qmom_order": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(_vp)]),
    "pbe_moc_state_create": (C.c_int, [_vp, _i64, _vp, C.POINTER(_vp)]),
    "pbe_moc_state_destroy": (C.c_int, [_vp]),
    "pbe_moc_state_upload": (C.c_int, [_vp, _i64, _vp]),
    "pbe_moc_state_download": (C.c_int, [_vp, _i64, _vp]),
    "pbe_moc_state_upload_range": (C.c_int, [_vp, _i64, _i64, _i64, _vp]),
    "pbe_moc_state_download_range": (C.c_int, [_vp, _i64, _i64, _i64, _vp]),
    "pbe_moc_state_d": (C.c_int, [_vp, _vp]),
    "pbe_moc_state_device": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_i64), C.POINTER(_vp)]),
    "pbe_moc_state_attach_mesh": (C.c_int, [_vp, _vp, C.c_int]),
    "pbe_moc_state_set_implicit": (C.c_int, [_vp, _vp]),
    "pbe_moc_state_solver_performance": (C.c_int, [_vp, _vp]),
    "pbe_moc_correct_resident": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, C.c_double, _vp,
                                           C.POINTER(MocDiag)]),
    "pbe_qmom_state_create": (C.c_int, [_vp, _i64, _vp, C.POINTER(_vp)]),
    "pbe_qmom_state_destroy": (C.c_int, [_vp]),
    "pbe_qmom_state_upload": (C.c_int, [_vp, _i64, _vp]),
    "pbe_qmom_state_download": (C.c_int, [_vp, _i64, _vp]),
    "pbe_qmom_state_d": (C.c_int, [_vp, _vp]),
    "pbe_qmom_state_device": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_i64), C.POINTER(_vp), C.POINTER(_vp)]),
    "pbe_qmom_state_attach_mesh": (C.c_int, [_vp, _vp, C.c_int]),
    "pbe_qmom_state_set_implicit": (C.c_int, [_vp, _vp]),
    "pbe_qmom_state_solver_performance": (C.c_int, [_vp, _vp]),
    "pbe_qmom_correct_resident": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, C.c_double, _vp,
                                            C.POINTER(QmomDiag)]),
    "pbe_comm_unique_id": (C.c_int, [_vp]),
    "pbe_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "pbe_comm_init_from": (C.c_int, [_vp, _vp]),
    "pbe_comm_destroy": (C.c_int, [_vp]),
    "pbe_comm_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pbe_reduce_diagnostics": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, _vp]),
    "pbe_reduce_diagnostics_host": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp]),
    "pbe_halo_exchange": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, _i64, _vp, _i64, _vp]),
    "pbe_host_register": (C.c_int, [_vp, _vp, _i64]),
    "pbe_host_unregister": (C.c_int, [_vp, _vp]),
    "pbe_moc_correct_host": (C.c_int, [_vp, _vp, C.c_int, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64,
                                       C.c_double, _vp]),
    "pbe_field_scrub_nan": (C.c_int, [_vp, _i64, _vp, _vp]),
    "pbe_qmom_clamp_d": (C.c_int, [_vp, C.c_int, _i64, _i64, _vp, _vp, _vp]),
    "pbe_mesh_set_coupled": (C.c_int, [_vp, _i64, _i64]),
    "pbe_gather_cells": (C.c_int, [_vp, C.c_int, _i64, _vp, _i64, _vp, _vp, _i64, _vp]),
    "pbe_transport_explicit_halo": (C.c_int, [_vp, C.c_int, C.c_int, _i64, _vp, _vp, _vp, _i64, _vp, _i64,
                                              _vp, C.c_double, _vp, _vp]),
    "pbe_transport_explicit_part": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _i64, _vp, _vp, _vp, _i64, _vp,
                                              _i64, _vp, C.c_double, _vp, _vp]),
    "pbe_eval_coalescence": (C.c_int, [_vp, C.POINTER(KernelSpec), _i64, _vp, _vp, _vp, _vp, _vp,
                                       _vp]),
    "pbe_eval_breakup": (C.c_int, [_vp, C.POINTER(KernelSpec), _i64, _vp, _vp, _vp, _vp]),
    "pbe_eval_daughter_beta": (C.c_int, [_vp, C.c_int, _i64, _vp, _vp, _vp]),
    "pbe_eval_daughter_moment": (C.c_int, [_vp, C.c_int, _i64, _vp, _vp, _vp]),
    "pbe_eval_integrate_exp": (C.c_int, [_vp, _i64, _vp, _vp, _vp]),
    "pbe_field_avg_max_min": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "pbe_device_alloc": (C.c_int, [_vp, _i64, C.POINTER(_vp)]),
    "pbe_device_free": (C.c_int, [_vp, _vp]),
    "pbe_host_alloc": (C.c_int, [_vp, _i64, C.POINTER(_vp)]),
    "pbe_host_free": (C.c_int, [_vp, _vp]),
    "pbe_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, _i64]),
    "pbe_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, _i64]),
    "pbe_memset": (C.c_int, [_vp, _vp, C.c_int, _i64]),
    "pbe_sync": (C.c_int, [_vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load the shared library; raise (never fall back) when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or make -C openpbe_b200/csrc).  openpbe_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the ABI lost a symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().pbe_last_error().decode(errors="replace")


def check(rc: int) -> None:
    if rc != PBE_OK:
        raise PbeError(rc, last_error())
