"""openpbe_b200 -- B200-native (sm_100a) population-balance source engine.

The product is the shared library `openpbe_b200/lib/libpbe_b200.so` (C ABI in
`include/pbe_b200.h`, CUDA sources in `openpbe_b200/csrc/`).  This package is the thin Python
plumbing around it used by tests/ and bench.py; importing it never falls back to a CPU path.
"""
from . import capi  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    # engine / synthetic import torch / numpy lazily so that `import openpbe_b200` stays cheap
    if name in ("Context", "MOC", "QMOM", "make_spec", "kernel_types"):
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)
