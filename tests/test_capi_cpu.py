"""CPU-side checks of the drop-in boundary (no compute calls: there is no GPU here).

 * libpbe_b200.so loads and exports every symbol include/pbe_b200.h declares;
 * run-time selection by TypeName behaves like the reference's tables (PBEMethod.C:81-108,
   coalescenceKernel.C:46-71): known names resolve, unknown names fail listing the valid ones;
 * without a CUDA device the library fails LOUDLY (PBE_E_CUDA) -- there is no CPU fallback;
 * the FP64 device math (pbe_math.cuh, __host__ __device__) agrees with libm when compiled by g++;
 * the product package never imports the oracle.
"""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pbe_b200.h")


@pytest.fixture(scope="module")
def capi():
    from openpbe_b200 import capi as m

    if not os.path.exists(m.LIB_PATH):
        import __graft_entry__ as g

        g.build()
    return m


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pbe_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported(capi):
    names = declared_functions()
    assert len(names) >= 35
    L = C.CDLL(capi.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"declared in include/pbe_b200.h but not exported: {missing}"
    # and the Python binding covers exactly the declared surface
    assert sorted(capi.PROTOTYPES) == names


def test_no_unexpected_cpu_symbols(capi):
    """The library must not carry the oracle (no orc_* symbols, no link to libpbe_oracle)."""
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in out
    ldd = subprocess.run(["ldd", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in ldd


def test_version_and_typenames(capi):
    L = capi.lib()
    assert L.pbe_version() == 100
    v = C.c_int(-1)
    for fn, table in ((L.pbe_coalescence_type, {"noCoalescence": 0, "constantCoalescence": 1,
                                                "CoulaloglouTavlaridesC": 2}),
                      (L.pbe_breakup_type, {"noBreakup": 0, "binaryBreakup": 1, "CoulaloglouTavlarides": 2}),
                      (L.pbe_daughter_type, {"uniformBinaryBreakup": 0})):
        for name, val in table.items():
            assert fn(name.encode(), C.byref(v)) == capi.PBE_OK
            assert v.value == val


def test_unknown_typename_lists_valid_ones(capi):
    L = capi.lib()
    v = C.c_int(-1)
    rc = L.pbe_coalescence_type(b"Madsen", C.byref(v))  # dead kernel in the reference: not selectable
    assert rc == capi.PBE_E_UNKNOWN_TYPE
    msg = capi.last_error()
    assert "Unknown coalescenceKernel type Madsen" in msg
    for valid in ("noCoalescence", "constantCoalescence", "CoulaloglouTavlaridesC"):
        assert valid in msg
    rc = L.pbe_breakup_type(b"LoAndRao", C.byref(v))
    assert rc == capi.PBE_E_UNKNOWN_TYPE and "binaryBreakup" in capi.last_error()
    assert L.pbe_daughter_type(None, C.byref(v)) == capi.PBE_E_INVALID


def test_fails_loudly_without_gpu(capi):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the no-device path is covered on the CPU box")
    h = C.c_void_p()
    rc = capi.lib().pbe_ctx_create(0, C.byref(h))
    assert rc == capi.PBE_E_CUDA
    assert not h
    assert "no CPU path" in capi.last_error()
    from openpbe_b200 import engine

    with pytest.raises(capi.PbeError):
        engine.Context(0)


def test_null_handles_are_rejected(capi):
    L = capi.lib()
    assert L.pbe_moc_source(None, 0, 0, None, None, None, None, 1.0, None, None) == capi.PBE_E_INVALID
    assert L.pbe_qmom_source(None, 0, 0, None, None, None, None, None, None) == capi.PBE_E_INVALID
    assert L.pbe_moc_create(None, 10, 1.0, None, 1.0, None) == capi.PBE_E_INVALID
    assert L.pbe_ctx_destroy(None) == capi.PBE_OK
    assert L.pbe_ctx_launch_count(None) == 0


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "openpbe_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn)).read()
                bad = re.findall(r"(?:import\s+oracle|from\s+oracle|oracle/|oracle\.|pbe_oracle|\borc_)", txt)
                assert not bad, f"{fn} reaches into oracle/: {bad}"


# ---- device math compiled for the host -------------------------------------------------------
HOST_MATH_SRC = r"""
#include "pbe_math.cuh"
#include <stdio.h>
#include <stdlib.h>
int main(int argc, char** argv) {
    long n = atol(argv[1]);
    srand48(12345);
    double worst = 0.0;
    for (long i = 0; i < n; ++i) {
        double u = drand48();
        double x = -pow(10.0, -9.0 + 11.85 * u);   /* -1e-9 .. -708 */
        if (x < -708.0) continue;
        double a = pbe::exp_neg(x);
        long double b = expl((long double)x);
        double rel = (double)fabsl(((long double)a - b) / b);
        if (rel > worst) worst = rel;
    }
    printf("%.17g\n", worst);
    printf("%.17g %.17g %.17g %.17g %.17g\n", pbe::exp_neg(0.0), pbe::exp_neg(-745.0), pbe::exp_neg(-1e300),
           pbe::exp_neg(-INFINITY), pbe::exp_neg(NAN));
    printf("%.17g %.17g %.17g %.17g\n", pbe::pow_third(27.0), pbe::pow_third(-8.0), pbe::pow_third(0.0),
           pbe::pow_int(0.0, 0));
    return 0;
}
"""


def test_device_math_on_host(tmp_path):
    src = tmp_path / "m.cpp"
    src.write_text(HOST_MATH_SRC)
    exe = tmp_path / "m"
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-I", os.path.join(ROOT, "openpbe_b200", "csrc"),
                           str(src), "-o", str(exe), "-lm"])
    out = subprocess.run([str(exe), "2000000"], capture_output=True, text=True, check=True).stdout.split("\n")
    worst = float(out[0])
    assert worst < 2.0 * 2.0 ** -53, f"exp_neg worst relative error {worst}"  # < 1 ulp
    e0, e745, ebig, einf, enan = (float(v) for v in out[1].split())
    assert e0 == 1.0 and e745 == 0.0 and ebig == 0.0 and einf == 0.0 and np.isnan(enan)
    c27, cneg, c0, p00 = (float(v) for v in out[2].split())
    assert c27 == 3.0 and np.isnan(cneg) and c0 == 0.0 and p00 == 1.0


# ---- synthetic workloads ---------------------------------------------------------------------
def test_synthetic_workloads_are_seeded_and_shardable():
    from openpbe_b200 import synthetic

    w = synthetic.workload("moc100_ct", 512)
    a = synthetic.moc_fields(w)
    b = synthetic.moc_fields(w)
    assert np.array_equal(a["n"], b["n"]) and np.array_equal(a["eps"], b["eps"])
    assert a["n"].shape == (100, 512) and np.all(a["n"] >= 0)
    assert 1e-3 <= a["eps"].min() and a["eps"].max() <= 1.0
    # volume fraction carried by the classes is alpha within the jitter band [0.5, 2]
    xi = w.xi1 * np.arange(1, 101)
    frac = (xi[:, None] * a["n"]).sum(0) / a["vcell"] / a["alpha"]
    assert np.all((frac > 0.5) & (frac < 2.0))
    # a shard is generated from its global offset, independent of the other shards
    s1 = synthetic.moc_fields(w, offset=512)
    assert not np.array_equal(s1["n"], a["n"])
    q = synthetic.qmom_fields(synthetic.workload("qmom3_ct", 4000))
    assert q["m"].shape == (6, 4000) and np.all(q["m"] > 0)
    deg = np.isclose(q["nodes"][1] / q["nodes"][0], 1 + 1e-6, rtol=1e-9)
    assert 10 <= deg.sum() <= 100  # about 1 % near-degenerate cells
    for name in ("moc30_ct", "qmom3_ct", "moc100_ct", "moc100_shard", "replay_moc", "replay_qmom"):
        assert synthetic.workload(name).ncells >= 100_000
