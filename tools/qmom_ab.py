"""A/B timing of the fused QMOM step (no S store) over library builds and grid sizes.
    python tools/qmom_ab.py            # spawns one process per library under openpbe_b200/lib/ab + the default
    python tools/qmom_ab.py --one LIB  # worker
"""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def worker(libpath):
    import numpy as np, torch
    from openpbe_b200 import capi
    capi.LIB_PATH = libpath
    from openpbe_b200 import engine, synthetic
    ctx = engine.Context(0)
    dev = torch.device("cuda:0")
    nc = 10_000_000
    w = synthetic.workload("qmom3_ct", nc)
    f = synthetic.qmom_fields(w, ncells=1_000_000)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    m0 = T(np.tile(f["m"], (1, 10))); e = T(np.tile(f["eps"], 10)); r = T(np.tile(f["rho"], 10)); v = T(np.tile(f["nu"], 10))
    q = engine.QMOM(ctx, 3, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    d = torch.empty(nc, dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    m = m0.clone()
    ref = None
    for waves in (0, 1, 2, 3, 4, 6, 8):
        os.environ["PBE_QMOM_WAVES"] = str(waves)
        ts = []
        for it in range(7):
            m.copy_(m0); flush.zero_()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(); q.step(m, e, r, v, w.dt, d=d); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts = sorted(ts[2:])
        print(json.dumps({"lib": os.path.basename(libpath), "waves": waves, "ms_median": ts[len(ts)//2], "ms_min": ts[0],
                          "checksum": float(m.sum().item()), "d_sum": float(d.sum().item())}), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        worker(sys.argv[2])
    else:
        libs = [os.path.join(ROOT, "openpbe_b200", "lib", "libpbe_b200.so")] + sorted(glob.glob(os.path.join(ROOT, "openpbe_b200", "lib", "ab", "lib*.so")))
        for lib in libs:
            subprocess.call([sys.executable, os.path.abspath(__file__), "--one", lib])
