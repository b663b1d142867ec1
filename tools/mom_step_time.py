"""Times pbe_mom_step (methodOfMoments) on the device: CUDA events on torch's
current stream (the stream the launch is given), 3 warm-ups, moments restored before every step.

    python tools/mom_step_time.py [ncells] [panels]
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openpbe_b200 import engine  # noqa: E402


def main():
    nc = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    panels = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(20240611)
    m0 = rng.uniform(0.5, 3.0, nc)
    k = rng.uniform(1.5, 4.0, nc)
    th = rng.uniform(0.15, 0.6, nc)
    m = torch.from_numpy(np.stack([m0, m0 * k * th, m0 * k * (k + 1) * th ** 2])).cuda()
    eps = torch.from_numpy(rng.uniform(0.5, 2.0, nc)).cuda()
    rho = torch.from_numpy(rng.uniform(0.8, 1.2, nc)).cuda()
    nu = torch.from_numpy(rng.uniform(0.5, 1.5, nc)).cuda()
    ctx = engine.Context(0)
    out = []
    CT = (1.0, 1.0, 0.0, 1.0)
    for name, spec in (("ct_ct", engine.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)),
                       ("const_binary", engine.make_spec("constantCoalescence", (0.1,), "binaryBreakup", ()))):
        mom = engine.MOM(ctx, spec, panels)
        work = m.clone()
        d = torch.empty(nc, dtype=torch.float64, device="cuda")
        ms = []
        for it in range(6):
            work.copy_(m)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            mom.step(work, eps, rho, nu, 1e-3, d=d)
            e1.record()
            torch.cuda.synchronize()
            if it >= 3:
                ms.append(e0.elapsed_time(e1))
        nq = 15 * ((panels or 12) + 20)
        t = float(np.mean(ms))
        out.append({"kernels": name, "cells": nc, "nodes": nq, "ms_per_step": t,
                    "cell_updates_per_s": nc / t * 1e3,
                    "kernel_evals_per_s": nc * 2.0 * nq * nq / t * 1e3})
        mom.close()
    for o in out:
        print(json.dumps(o))


if __name__ == "__main__":
    main()
