#!/usr/bin/env python
"""Time pbe_moc_correct_resident / pbe_qmom_correct_resident (host phase fields in, d out) against
the resident device step, for several chunk counts.  Prints one JSON line per measurement.

    python tools/resident_time.py [--cells 16000000] [--chunks 1,4,8,12,16,24,32]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from openpbe_b200 import engine, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=float, default=16_000_000)
    ap.add_argument("--qcells", type=float, default=10_000_000)
    ap.add_argument("--chunks", default="1,4,8,12,16,24,32")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--workload", default="moc100_ct")
    args = ap.parse_args()
    cells = int(args.cells)
    ctx = engine.Context(0)
    dev = torch.device("cuda:0")
    w = synthetic.workload(args.workload)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    blk = 500_000
    st = engine.MOCState(moc, cells)
    n_dev, _ = st.device_fields()
    hs = {k: torch.empty(cells, dtype=torch.float64, pin_memory=True) for k in ("eps", "rho", "nu", "alpha", "vcell")}
    for c0 in range(0, cells, blk):
        nb = min(blk, cells - c0)
        f = synthetic.moc_fields(w, offset=c0, ncells=nb)
        n_dev[:, c0:c0 + nb].copy_(torch.from_numpy(f["n"]))
        for k in hs:
            hs[k][c0:c0 + nb].copy_(torch.from_numpy(f[k]))
    n0 = n_dev.clone()
    torch.cuda.synchronize()
    del n_dev
    st.close()
    d_h = torch.empty(cells, dtype=torch.float64, pin_memory=True)
    dv = {k: v.to(dev) for k, v in hs.items()}
    # device-resident step for reference
    n = n0.clone()
    dd = torch.empty(cells, dtype=torch.float64, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        moc.step(n, dv["eps"], dv["rho"], dv["nu"], w.dt, d=dd)
    e0.record()
    for _ in range(args.steps):
        moc.step(n, dv["eps"], dv["rho"], dv["nu"], w.dt, d=dd)
    e1.record()
    torch.cuda.synchronize()
    base = e0.elapsed_time(e1) / args.steps
    print(json.dumps({"what": "pbe_moc_step resident", "cells": cells, "ms": base}), flush=True)
    del n
    for variant in ("full", "nodiag", "pageable"):
        for ch in [int(x) for x in args.chunks.split(",")]:
            if variant != "full" and ch not in (1, 12):
                continue
            os.environ["PBE_RESIDENT_CHUNKS"] = str(ch)
            st = engine.MOCState(moc, cells, vcell=hs["vcell"])
            sn, _ = st.device_fields()
            sn.copy_(n0)
            torch.cuda.synchronize()
            src = hs if variant != "pageable" else {k: v.clone() for k, v in hs.items()}  # clone(): pageable
            dout = d_h if variant != "pageable" else torch.empty(cells, dtype=torch.float64)
            kw = dict(minmax=variant == "full", n_moments=6 if variant == "full" else 0)
            st.correct(src["eps"], src["rho"], src["nu"], src["alpha"], w.dt, d_out=dout, **kw)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                st.correct(src["eps"], src["rho"], src["nu"], src["alpha"], w.dt, d_out=dout, **kw)
            ms = (time.perf_counter() - t0) / args.steps * 1e3
            print(json.dumps({"what": f"pbe_moc_correct_resident {variant}", "chunks": ch, "cells": cells, "ms": ms,
                              "vs_device_step": ms / base}), flush=True)
            st.close()
    del n0, dv
    # ---- QMOM ---------------------------------------------------------------------------------------
    qc = int(args.qcells)
    wq = synthetic.workload("qmom3_ct")
    q = engine.QMOM(ctx, wq.N, engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs))
    qh = {k: torch.empty(qc, dtype=torch.float64, pin_memory=True) for k in ("eps", "rho", "nu")}
    m0 = torch.empty((2 * wq.N, qc), dtype=torch.float64)
    for c0 in range(0, qc, 1_000_000):
        nb = min(1_000_000, qc - c0)
        f = synthetic.qmom_fields(wq, offset=c0, ncells=nb)
        m0[:, c0:c0 + nb] = torch.from_numpy(f["m"])
        for k in qh:
            qh[k][c0:c0 + nb] = torch.from_numpy(f[k])
    dq = torch.empty(qc, dtype=torch.float64, pin_memory=True)
    for ch in (1, 4, 12, 24):
        os.environ["PBE_RESIDENT_CHUNKS"] = str(ch)
        qs = engine.QMOMState(q, qc, m0=m0)
        qs.correct(qh["eps"], qh["rho"], qh["nu"], wq.dt, d_out=dq)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            qs.correct(qh["eps"], qh["rho"], qh["nu"], wq.dt, d_out=dq)
        ms = (time.perf_counter() - t0) / args.steps * 1e3
        print(json.dumps({"what": "pbe_qmom_correct_resident", "chunks": ch, "cells": qc, "ms": ms,
                          "pcie_gbs": qc * 32 / ms / 1e6}), flush=True)
        qs.close()


if __name__ == "__main__":
    main()
