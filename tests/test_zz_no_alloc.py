"""pbe_moc_class_minmax_noalloc: the allocation-free variant of pbe_moc_class_minmax (no counterpart in the
reference, include/pbe_b200.h).  Collected last on purpose.
Called through the C ABI, compared with the existing entry point --
min / max do not depend on the order of the partials, so the tolerance is exact equality -- on two calls
in a row (the calls share one static scratch array that keeps the previous call's contents)."""
import ctypes as C

import numpy as np
import pytest

gpu = pytest.mark.gpu


def _minmax(lib, fn, moc, n, stream):
    import torch
    N, nc = n.shape
    lo = torch.full((N,), float("nan"), dtype=torch.float64, device=n.device)
    hi = torch.full((N,), float("nan"), dtype=torch.float64, device=n.device)
    rc = getattr(lib, fn)(moc._h, nc, n.stride(0), C.c_void_p(n.data_ptr()), C.c_void_p(lo.data_ptr()),
                          C.c_void_p(hi.data_ptr()), stream)
    return rc, lo, hi


@gpu
def test_noalloc_class_minmax_matches_allocating_entry_point():
    import torch
    from openpbe_b200 import capi, engine
    lib = capi.lib()
    ctx = engine.Context(0)
    dev = torch.device("cuda:0")
    spec = engine.make_spec("constantCoalescence", (1.0,), "noBreakup", ())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    # two calls in a row on DIFFERENT shapes: the second finds the first one's partials in the scratch
    # (more partials, larger values) and must not see them
    cases = [(100, 700_001, 1e9), (30, 5_003, 1.0), (100, 700_001, 1e9), (1, 1, 3.0)]
    mocs = {N: engine.MOC(ctx, N, 1e-11, spec) for N in {c[0] for c in cases}}
    for N, nc, scale in cases:
        n = (torch.rand(N, nc, dtype=torch.float64, device=dev, generator=gen) - 0.25) * scale
        rc0, lo0, hi0 = _minmax(lib, "pbe_moc_class_minmax", mocs[N], n, stream)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc1, lo1, hi1 = _minmax(lib, "pbe_moc_class_minmax_noalloc", mocs[N], n, stream)
        rc2, lo2, hi2 = _minmax(lib, "pbe_moc_class_minmax_noalloc", mocs[N], n, stream)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0 and rc2 == 0, capi.last_error()
        for lo, hi in ((lo1, hi1), (lo2, hi2)):
            assert torch.equal(lo, lo0) and torch.equal(hi, hi0)
        assert torch.equal(lo0, n.min(dim=1).values) and torch.equal(hi0, n.max(dim=1).values)
        # lo / hi tensors come from torch's caching allocator; the entry point itself takes nothing from the device
        assert torch.cuda.mem_get_info(dev)[0] >= free_before - (4 << 20)
    # empty field: identities of min / max, as the existing entry point
    e = torch.empty(30, 0, dtype=torch.float64, device=dev)
    rc0, lo0, hi0 = _minmax(lib, "pbe_moc_class_minmax", mocs[30], e, stream)
    rc1, lo1, hi1 = _minmax(lib, "pbe_moc_class_minmax_noalloc", mocs[30], e, stream)
    torch.cuda.synchronize()
    assert rc0 == 0 and rc1 == 0 and torch.equal(lo1, lo0) and torch.equal(hi1, hi0)
    for m in mocs.values():
        m.close()


@gpu
def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from openpbe_b200 import capi, engine
    lib = capi.lib()
    ctx = engine.Context(0)
    dev = torch.device("cuda:0")
    spec = engine.make_spec("noCoalescence", (), "noBreakup", ())
    moc = engine.MOC(ctx, 129, 1e-11, spec)  # PBE_NOALLOC_MAX_CLASSES = 128
    n = torch.rand(129, 1000, dtype=torch.float64, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rc, _, _ = _minmax(lib, "pbe_moc_class_minmax_noalloc", moc, n, stream)
    assert rc == capi.PBE_E_UNSUPPORTED and "exceeds the static scratch's limit 128" in capi.last_error()
    rc, lo, hi = _minmax(lib, "pbe_moc_class_minmax", moc, n, stream)  # the allocating one takes it
    torch.cuda.synchronize()
    assert rc == 0 and torch.equal(lo, n.min(dim=1).values)
    assert lib.pbe_moc_class_minmax_noalloc(None, 0, 0, None, None, None, None) == capi.PBE_E_INVALID
    moc.close()


def test_noalloc_symbol_and_limits_declared():
    """CPU: the library exports the entry point and the header states the limits."""
    import os
    from openpbe_b200 import capi
    assert hasattr(capi.lib(), "pbe_moc_class_minmax_noalloc")
    hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "pbe_b200.h")).read()
    assert "#define PBE_NOALLOC_MAX_CLASSES 128" in hdr and "#define PBE_NOALLOC_MAX_PARTS 512" in hdr
    assert "NO COUNTERPART IN THE REFERENCE" in hdr
