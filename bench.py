#!/usr/bin/env python
"""bench.py -- PBE source throughput of libpbe_b200.so on N B200s (one process per GPU).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU implementation

Metric (BASELINE.json): MOC cell-class-updates/s at 100 classes.  One *step* = one pass of the
hot path over this rank's cells: `pbe_moc_step` = breakup-rate evaluation + coalescence
birth/death double sums + breakup redistribution + limiter + zero-flux explicit update in place
(MOC::correct() with phi = 0, src/PBESystem/MOC.C:159-184, 259-272) with the fused global-moment
partial sums of BASELINE configs[3]; for N > 1 the 6 partial moments are all-reduced over NCCL on
a side stream (the only cross-rank traffic: cells are independent).

  * N = 1 workload = BASELINE configs[2]: 100 classes, 1.6e7 cells, Coulaloglou-Tavlarides
    coalescence + breakup with the c01 validation constants, synthetic eps / alpha fields.
  * N > 1: 1.6e7 cells PER GPU (weak scaling; N = 8 is configs[3]'s 1.28e8 cells); `strong` adds
    configs[3] as written: 1.28e8 cells in total, split over the N GPUs.
  * `value`  : whole-job cell-class-updates/s, fields resident in HBM, CUDA-event time of exactly
               K steps between barriers, max over ranks.  Inputs (12.8 GB / rank) exceed L2.
  * `e2e`    : the same metric through the call the solver makes once per time step,
               `pbe_moc_correct_resident`: the class fields are owned by the method on the device
               (as the reference's MOC owns classNumberDensity_, MOC.H:66), this step's phase fields
               (eps, rho, nu, alpha: 32 B/cell, pinned host memory) go in and d comes out (8 B/cell),
               copies inside the timed region, with the per-class minima, the continuity error and
               the 6 global moments MOC::correct() prints.  `e2e.roundtrip` keeps round 1's number:
               all class fields over the bus both ways every step (`pbe_moc_source_host`).
  * `roofline`: the MOC kernel is FP64-pipe bound at 100 classes (66 flop/B); `roofline` reports
               the binding FP64 roof (peak = DFMA-chain micro-kernel measured in the same run --
               MEASURED_PEAKS.json has no FP64 figure) and `roofline.hbm` the HBM view against
               MEASURED_PEAKS.json.  `roofline.legs` repeats the roofline of every other leg.
  * `cpu_baseline`: the reference's own MOC.C compiled unmodified against an OpenFOAM API shim
               (oracle/_ref, kind "reference") when it was built, else oracle/pbe_oracle.c (kind
               "port"), on a bounded slice of the same workload, all host threads.
               `cpu_baseline.legs` holds the baselines of the other legs.
  * `config1`: BASELINE configs[0] (MOC 30 classes, 1e5 cells, benchmark constants) with the FULL
               1e5-cell CPU baseline.
  * `qmom`   : the second half of the metric (QMOM 6 moments, BASELINE configs[1], 1e7 cells / GPU)
               with roofline = max(HBM, FP64) and a reference-shaped CPU baseline (one inversion per
               moment index, QMOM.C:231).
  * `replay` : BASELINE configs[4]: PBEMethod::correct() per time step on 4e6 cells, MOC (50 classes,
               c01) and QMOM (order 3) side by side through the resident correct(), ms/step each.
  * `transport`, `mom`: SURVEY 8 rows f2 / f3.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "moc100_cell_class_updates_per_s"
UNIT = "cell-class-updates/s"

# ---- algorithmic work per cell ----------------------------
# (1) SURVEY.md 8(d) convention -- what `roofline.achieved` uses: add/mul = 1, FMA = 2, div/sqrt = 1,
#     one FP64 exp = 32; P(N) = unordered class pairs with i + j <= N - 2:
#       MOC + CT coalescence + CT breakup: 40 + 41 P(N) + 40 N  flops / cell  (106 540 at N = 100)
#       MOC + constant / table kernels:     9 P(N) + 8 N
# (2) SASS-counted FP64-pipe instructions of the kernel that runs (pbe_moc_tmem_kernel.cuh, cuobjdump
#     of the tile loop and the epilogue).  The FP64 pipe retires one warp instruction per scheduler
#     every 2 cycles whatever the opcode, so instructions x 2 / DFMA-peak is the pipe utilisation
#     (`roofline.fp64_pipe_frac`).
FP64_INSTR_PER_PAIR_CT = 14.3   # 13.25 per table slot x 2704 slots / 2500 valid pairs at N = 100 (8 % zero padding)
FP64_INSTR_PER_PAIR_TAB = 4.0
FP64_INSTR_PER_CLASS_CT = 42.3  # ncu r01w: 39 975 FP64 warp instructions per cell in total at N = 100
FP64_INSTR_PER_CLASS_TAB = 30.0
FLOPS_QMOM3_CT = 1300.0
FLOPS_QMOM3_CONST = 600.0


def n_pairs(N: int) -> int:
    """unordered pairs (i <= j) with i + j <= N - 2: the kernel evaluations one cell needs"""
    return sum(max(0, N - 1 - 2 * i) for i in range(N // 2 + 1))


def moc_flops_per_cell(w) -> float:
    """SURVEY.md 8(d) algorithmic flops per cell (exp = 32 flops)."""
    ct_c = w.coalescence == "CoulaloglouTavlaridesC"
    ct_b = w.breakup == "CoulaloglouTavlarides"
    pairs = 0 if w.coalescence == "noCoalescence" else n_pairs(w.N)
    if ct_c or ct_b:
        return 40.0 + (41.0 if ct_c else 9.0) * pairs + (40.0 if ct_b else 8.0) * w.N
    return 9.0 * pairs + 8.0 * w.N


def moc_fp64_instr_per_cell(w) -> float:
    """SASS-counted FP64-pipe instructions per cell of the tensor-memory kernel."""
    ct_c = w.coalescence == "CoulaloglouTavlaridesC"
    ct_b = w.breakup == "CoulaloglouTavlarides"
    pairs = 0 if w.coalescence == "noCoalescence" else n_pairs(w.N)
    return pairs * (FP64_INSTR_PER_PAIR_CT if ct_c else FP64_INSTR_PER_PAIR_TAB) + w.N * (
        FP64_INSTR_PER_CLASS_CT if ct_b else FP64_INSTR_PER_CLASS_TAB)


def moc_bytes_per_cell(N: int) -> float:
    return 8.0 * (2 * N + 3)  # read N n_i + eps, rho, nu; write N n_i (or S_i)


def load_traffic(key: str, cells: int):
    """dram bytes per launch from the committed ncu captures (profiles/r02_traffic.json, else r01),
    scaled by the cell count when the bench runs another size; None when there is no capture.
    This is a committed constant of an earlier `ncu --set full` run, NOT measured by this run."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                t = json.load(fh)[key]
            per_cell = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["cells"]
            return per_cell * cells, f"profiles/{name}"
        except Exception:
            continue
    return None, None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- clocks during the timed region --------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, pw, reasons = [], [], [], set()
        for ts, ln in self.lines:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                               f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no sample inside the timed region"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- synthetic fields, generated block-wise into (pinned) host memory ------------------------------
def fill_moc_host(w, ncells, cell_offset, n_host, eps_host, rho_host, nu_host, alpha_host, vcell_host,
                  block=250_000):
    """SURVEY.md 8(d) fields; block-wise so that each (seed, offset) block is reproducible."""
    from concurrent.futures import ThreadPoolExecutor

    from openpbe_b200 import synthetic

    def one(b0):
        nb = min(block, ncells - b0)
        f = synthetic.moc_fields(w, offset=cell_offset + b0, ncells=nb)
        n_host[:, b0:b0 + nb] = f["n"]
        eps_host[b0:b0 + nb] = f["eps"]
        rho_host[b0:b0 + nb] = f["rho"]
        nu_host[b0:b0 + nb] = f["nu"]
        alpha_host[b0:b0 + nb] = f["alpha"]
        vcell_host[b0:b0 + nb] = f["vcell"]

    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(one, range(0, ncells, block)))


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ==================================================================================================
# CPU arms: the reference's own sources (oracle/_ref) when they were built, else the oracle port
# ==================================================================================================
def host_threads() -> int:
    """Host threads this process may use.  Not omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to every rank, which would time the reference arm on one core."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_arm():
    """(kind, moc(w, fields, nthreads, nsteps, warmup) -> seconds per correct(), description)"""
    from oracle import cpu_oracle as orc

    orc.build()
    try:
        from oracle import ref_oracle as ref

        if ref.available():
            def run_ref(w, f, nthreads, nsteps=1, warmup=0):
                return ref.correct_timed("methodOfClasses", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, f["n"],
                                         f["eps"], f["rho"], f["nu"], w.dt, N=w.N, xi1=w.xi1, relax=w.relax,
                                         nthreads=nthreads, nsteps=nsteps, warmup=warmup)

            return "reference", run_ref, ("the reference's own MOC.C / CoulaloglouTavlarides.C / uniformBinaryBreakup.C "
                                          "compiled unmodified against the OpenFOAM API shim (oracle/_ref): "
                                          "MOC::correct() + d() in a time loop; one process per host thread on "
                                          "disjoint cell blocks, as decomposePar + mpirun runs it")
    except Exception:  # noqa: BLE001 -- no _ref on this box: fall back to the port, and say so in `kind`
        pass

    def run_port(w, f, nthreads, nsteps=1, warmup=0):
        sp = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
        for _ in range(warmup):
            orc.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt, nthreads=nthreads)
        t = time.perf_counter()
        for _ in range(nsteps):
            orc.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt, nthreads=nthreads)
        return (time.perf_counter() - t) / max(1, nsteps)

    return "port", run_port, ("oracle/pbe_oracle.c reference-shaped loops (MOC.C:92-157 order, pow/exp per ordered "
                              "pair), OpenMP over cells")


def cpu_moc_rate(w, sample_cells: int, nthreads: int, run, repeats: int = 1):
    """cell-class-updates/s of the CPU arm on a slice of the workload."""
    from openpbe_b200 import synthetic

    f = synthetic.moc_fields(w, ncells=sample_cells)
    best = min(run(w, f, nthreads) for _ in range(repeats))
    return sample_cells * w.N / best, best


def cpu_qmom_rate(wq, sample_cells: int, nthreads: int):
    """(cell-updates/s, seconds, kind, how) of the reference-shaped QMOM correct(): one wheeler_inversion per
    moment index (QMOM.C:231; benchmarks/moment_inversion.cpp:5-12)."""
    from openpbe_b200 import synthetic
    from oracle import cpu_oracle as orc

    f = synthetic.qmom_fields(wq, ncells=sample_cells)
    try:
        from oracle import ref_oracle as ref

        if ref.available():
            secs = ref.correct_timed("quadratureMethodOfMoments", wq.coalescence, wq.coal_coeffs, wq.breakup,
                                     wq.br_coeffs, f["m"], f["eps"], f["rho"], f["nu"], wq.dt, order=wq.N,
                                     nthreads=nthreads)
            return sample_cells / secs, secs, "reference", ("the reference's own QMOM.C + MomentInversion.C compiled "
                                                            "unmodified (oracle/_ref); Eigen is absent: its "
                                                            "SelfAdjointEigenSolver is the shim's Jacobi rotations")
    except Exception:  # noqa: BLE001
        pass
    sp = orc.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs)
    t = time.perf_counter()
    orc.qmom_source(sp, wq.N, f["m"], f["eps"], f["rho"], f["nu"], reference_shaped=True, nthreads=nthreads)
    secs = time.perf_counter() - t
    return sample_cells / secs, secs, "port", "oracle/pbe_oracle.c, reference-shaped (MomentInversion.C needs Eigen)"


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    from openpbe_b200 import synthetic

    kind, run, how = cpu_arm()
    w = synthetic.workload(args.workload)
    threads = host_threads()
    sample = args.cpu_sample_cells
    f = synthetic.moc_fields(w, ncells=sample)
    # one time loop: `warmup` untimed correct() calls, then `steps` timed ones (fields evolve, as in PBEFoam)
    dt = run(w, f, threads, nsteps=max(1, args.steps), warmup=args.warmup)
    value = sample * w.N / dt
    desc = f"{sample} cells of {w.name} (N={w.N}) per step; {how}; {threads} host threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{w.name}: MOC {w.N} classes, CT coalescence + CT breakup (c01 constants)",
                   "cells_per_step": sample, "note": "CPU rate is per-cell; the sample bounds run time"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ==================================================================================================
# this repo's arm
# ==================================================================================================
class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from openpbe_b200 import engine

        self.args = args
        self.torch, self.dist = torch, dist
        self.rank, self.world, self.local = dist_env()
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- libpbe_b200 has no CPU path")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            # NCCL's own log (NCCL_DEBUG) is left alone: the JSON line is printed last
            dist.init_process_group("nccl", device_id=self.dev)
            t = torch.zeros(1, device=self.dev)
            dist.all_reduce(t)  # creates the communicator now, outside every timed region
            torch.cuda.synchronize()
        self.ctx = engine.Context(self.local)
        self.info = self.ctx.info()
        self.hbm_peak, self.hbm_src = load_peaks()
        self.fp64_tf, _ = self.ctx.fp64_peak_tflops()
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.flush_buf = None

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def allmax(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def flush_l2(self):
        if self.flush_buf is None:
            self.flush_buf = self.torch.empty(256 << 20, dtype=self.torch.uint8, device=self.dev)  # > 126 MB L2
        self.flush_buf.zero_()

    def pinned(self, a):
        return self.torch.from_numpy(np.ascontiguousarray(a)).pin_memory()


def leg_main(B: Bench):
    """configs[2] / configs[3] (weak): the headline value, its roofline, e2e and the strong-scaling line"""
    torch, dist, args = B.torch, B.dist, B.args
    from openpbe_b200 import engine, synthetic

    rank, world, dev, ctx = B.rank, B.world, B.dev, B.ctx
    w = synthetic.workload(args.workload)
    cells = int(args.cells) if args.cells else w.ncells  # per GPU (weak scaling)
    N, nmom = w.N, 6
    cell0 = rank * cells

    # ---- inputs: generated chunk-wise on the host; resident copies in HBM; the per-cell phase fields
    # also stay in pinned host memory (what the solver hands to correct() every step) ---------------
    t_gen = time.perf_counter()
    host = {k: torch.empty(cells, dtype=torch.float64, pin_memory=True) for k in ("eps", "rho", "nu", "alpha", "vcell")}
    rt_cells = 0 if args.no_e2e else min(cells, int(args.roundtrip_cells))
    n_host = torch.empty((N, max(1, rt_cells)), dtype=torch.float64, pin_memory=True)
    S_host = torch.empty((N, max(1, rt_cells)), dtype=torch.float64, pin_memory=True)
    n0 = torch.empty((N, cells), dtype=torch.float64, device=dev)
    chunk = min(cells, 2_000_000)
    st_n = np.empty((N, chunk))
    st_s = [np.empty(chunk) for _ in range(5)]
    for c0 in range(0, cells, chunk):
        nb = min(chunk, cells - c0)
        fill_moc_host(w, nb, cell0 + c0, st_n, *st_s)
        n0[:, c0:c0 + nb].copy_(torch.from_numpy(st_n[:, :nb]))
        for key, src in zip(("eps", "rho", "nu", "alpha", "vcell"), st_s):
            host[key][c0:c0 + nb].copy_(torch.from_numpy(src[:nb]))
        if c0 < rt_cells:
            k = min(nb, rt_cells - c0)
            n_host[:, c0:c0 + k].copy_(torch.from_numpy(st_n[:, :k]))
    del st_n, st_s
    eps, rho, nu, alpha, vcell = (host[k].to(dev) for k in ("eps", "rho", "nu", "alpha", "vcell"))
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t_gen

    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, N, w.xi1, spec, w.relax)
    n = n0.clone()
    moments = torch.zeros(nmom, dtype=torch.float64, device=dev)
    side = torch.cuda.Stream(device=dev)
    red_bufs = []

    def make_step(nn, e_, r_, v_, vc_):
        def step():
            moments.zero_()
            moc.step(nn, e_, r_, v_, w.dt, vcell=vc_, moments=moments)
            if world > 1:
                # global moments sum_c V_c sum_i xi_i^k n_i: 48 B, latency-bound, off the kernel stream
                buf = moments.clone()
                side.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side):
                    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
                red_bufs.append(buf)
        return step

    step = make_step(n, eps, rho, nu, vcell)
    warmup = max(args.warmup, 3)  # timing rule: at least 3 untimed steps
    for _ in range(warmup):
        step()
    red_bufs.clear()
    sampler = ClockSampler(B.local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    B.barrier()
    launches0 = ctx.launch_count()
    tw0 = time.perf_counter()
    B.e0.record()
    for _ in range(args.steps):
        step()
    B.e1.record()
    torch.cuda.current_stream(dev).wait_stream(side)
    B.barrier()
    tw1 = time.perf_counter()
    launches = ctx.launch_count() - launches0
    ms_local = B.e0.elapsed_time(B.e1)
    ms = B.allmax(ms_local)
    clocks = sampler.stop(tw0, tw1) if rank == 0 else None
    ms_per_step = ms / args.steps
    value = world * cells * N / (ms_per_step * 1e-3)
    gmom = red_bufs[-1].cpu().tolist() if red_bufs else moments.cpu().tolist()
    red_bufs.clear()
    # realisability after the replayed steps: n >= 0 everywhere (MOC limiter, MOC.C:87-91)
    n_min = float(n.min().item())

    # ---- roofline of the dominant (only) kernel: one launch per step ---------------------------
    kern_ms = ms_local / args.steps
    flops_cell = moc_flops_per_cell(w)
    bytes_cell = moc_bytes_per_cell(N) + 8.0  # + vcell for the moment sums
    tf = cells * flops_cell / (kern_ms * 1e-3) / 1e12
    gbs = cells * bytes_cell / (kern_ms * 1e-3) / 1e9
    fp64_nominal = B.info["sm_count"] * 64 * 2 * 1.965e9 / 1e12
    kinfo = moc.kernel()
    instr_cell = moc_fp64_instr_per_cell(w)
    pipe_frac = cells * instr_cell * 2.0 / (kern_ms * 1e-3) / 1e12 / B.fp64_tf
    traffic, traffic_src = (load_traffic("moc_tmem_kernel", cells) if (kinfo["kernel"] == "tmem" and N == 100)
                            else (None, None))
    roofline = {
        "bound": "fp64", "achieved": tf, "peak": B.fp64_tf, "unit": "TFLOP/s", "frac": tf / B.fp64_tf,
        "traffic": traffic,
        "traffic_source": f"committed constant, not measured by this run: ncu --set full dram__bytes_read.sum + "
                          f"dram__bytes_write.sum of {traffic_src} (captured at 1.6e7 cells; scaled per cell)",
        "algorithmic_bytes": cells * bytes_cell, "kernel_ms": kern_ms,
        "kernel": ("moc_tmem_kernel<CT,CT>" if kinfo["kernel"] == "tmem" else "moc_source_kernel<CT,CT,128>")
                  + f" ({kinfo['cells_per_cta']} cells/CTA, {kinfo['ctas_per_sm']} CTA/SM)",
        "peak_source": "measured in this run: register-resident DFMA chains (pbe_measure_fp64_peak); "
                       f"nominal {fp64_nominal:.1f} TFLOP/s at 1965 MHz; MEASURED_PEAKS.json has no FP64 figure",
        "flops_per_cell": flops_cell,
        "flops_convention": "SURVEY.md 8(d): add/mul=1, FMA=2, exp=32: 40 + 41 P(N) + 40 N per cell",
        "fp64_pipe_frac": pipe_frac, "fp64_pipe_instr_per_cell": instr_cell,
        "fp64_pipe_note": "SASS-counted FP64-pipe instructions x 2 / DFMA peak = share of FP64 issue slots used",
        "hbm": {"bound": "hbm", "achieved": gbs, "peak": B.hbm_peak, "unit": "GB/s", "frac": gbs / B.hbm_peak,
                "bytes_per_cell": bytes_cell, "peak_source": B.hbm_src},
        "legs": {},
    }
    del n

    # ---- strong scaling: configs[3] as written (1.28e8 cells in total) over the N GPUs ---------------
    strong = None
    if world > 1 and not args.no_strong:
        total = int(args.strong_cells)
        sc = total // world
        if sc == cells:
            strong = {"global_cells": total, "cells_per_gpu": sc, "ms_per_step": ms_per_step, "value": value,
                      "note": "identical to the weak-scaling run at this N (1.6e7 cells per GPU)"}
        else:
            # the per-GPU block is the weak-scaling block repeated (device copies): same per-cell work
            ns = torch.empty((N, sc), dtype=torch.float64, device=dev)
            small = [torch.empty(sc, dtype=torch.float64, device=dev) for _ in range(4)]
            for c0 in range(0, sc, cells):
                k = min(cells, sc - c0)
                ns[:, c0:c0 + k].copy_(n0[:, :k])
                for dst, src in zip(small, (eps, rho, nu, vcell)):
                    dst[c0:c0 + k].copy_(src[:k])
            sstep = make_step(ns, *small)
            for _ in range(2):
                sstep()
            red_bufs.clear()
            B.barrier()
            ssteps = max(3, min(args.steps, 5))
            B.e0.record()
            for _ in range(ssteps):
                sstep()
            B.e1.record()
            torch.cuda.current_stream(dev).wait_stream(side)
            B.barrier()
            sms = B.allmax(B.e0.elapsed_time(B.e1)) / ssteps
            red_bufs.clear()
            strong = {"global_cells": sc * world, "cells_per_gpu": sc, "ms_per_step": sms, "steps": ssteps,
                      "value": sc * world * N / (sms * 1e-3), "unit": UNIT, "scaling": "strong",
                      "bytes_per_gpu": sc * N * 8,
                      "note": "configs[3]: 1.28e8 cells block-sharded over the GPUs, in-place step + NCCL all-reduce "
                              "of 6 global moments; the per-GPU block repeats the 1.6e7-cell synthetic block"}
            del ns, small

    # ---- e2e: the call the solver makes once per time step ---------------------------------------
    e2e = None
    if not args.no_e2e:
        st = engine.MOCState(moc, cells, vcell=host["vcell"])
        sn, _ = st.device_fields()
        sn.copy_(n0)
        torch.cuda.synchronize()
        d_host = torch.empty(cells, dtype=torch.float64, pin_memory=True)
        red = torch.empty(N + nmom, dtype=torch.float64, device=dev)

        def estep():
            d, diag = st.correct(host["eps"], host["rho"], host["nu"], host["alpha"], w.dt, d_out=d_host,
                                 minmax=True, n_moments=nmom)
            if world > 1:  # the global min(n_k) and moments MOC::correct() prints (MOC.C:181-183)
                red[:N].copy_(torch.from_numpy(diag["class_min"]))
                red[N:].copy_(torch.tensor(diag["moments"], dtype=torch.float64))
                dist.all_reduce(red[:N], op=dist.ReduceOp.MIN)
                dist.all_reduce(red[N:], op=dist.ReduceOp.SUM)
            return diag

        # the chunked pipeline must give exactly what one resident step gives
        estep()
        chk = min(cells, 4096)
        nchk = n0[:, :chk].clone()
        moc.step(nchk, eps[:chk].contiguous(), rho[:chk].contiguous(), nu[:chk].contiguous(), w.dt)
        if not torch.equal(nchk, sn[:, :chk]):
            raise SystemExit("bench.py: resident correct() and the device step disagree")
        for _ in range(2):
            estep()
        e2e_steps = max(3, min(args.steps, args.e2e_steps))
        B.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            diag = estep()
        torch.cuda.synchronize()
        t_e2e = B.allmax((time.perf_counter() - t0) / e2e_steps)
        B.barrier()
        h2d = 4 * cells * 8
        d2h = cells * 8 + (N * 4 + 9) * 8
        e2e = {"value": world * cells * N / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": t_e2e * 1e3, "steps": e2e_steps,
               "call": "pbe_moc_correct_resident: class fields owned on the device (MOC.H:66); per step pinned host "
                       "eps, rho, nu, alpha -> d, per-class min / max, continuity error, 6 global moments; "
                       "synchronous; chunked copy | kernel | copy pipeline"
                       + ("; NCCL all-reduce of the minima and moments" if world > 1 else ""),
               "pinned": True, "pcie_gbs": (h2d + d2h) / t_e2e / 1e9, "cells_per_gpu": cells,
               "kernel_ms_per_step": ms_per_step, "overhead_vs_resident_kernel": t_e2e * 1e3 / ms_per_step,
               "cont_err_max": diag["cont_err_max"], "legs": {}}
        del sn
        st.close()
        # round 1's e2e, kept beside it: every class field over the bus both ways, every step
        moc.source_host(n_host, host["eps"][:rt_cells], host["rho"][:rt_cells], host["nu"][:rt_cells], w.dt,
                        out=S_host)
        B.barrier()
        t0 = time.perf_counter()
        for _ in range(2):
            moc.source_host(n_host, host["eps"][:rt_cells], host["rho"][:rt_cells], host["nu"][:rt_cells], w.dt,
                            out=S_host)
        torch.cuda.synchronize()
        t_rt = B.allmax((time.perf_counter() - t0) / 2)
        B.barrier()
        e2e["roundtrip"] = {"value": world * rt_cells * N / t_rt, "unit": UNIT, "ms_per_step": t_rt * 1e3,
                            "cells_per_gpu": rt_cells, "h2d_bytes_per_step": (N + 3) * rt_cells * 8,
                            "d2h_bytes_per_step": N * rt_cells * 8,
                            "pcie_gbs": (2 * N + 3) * rt_cells * 8 / t_rt / 1e9,
                            "call": "pbe_moc_source_host (pinned host n, eps, rho, nu -> S): all class fields over "
                                    "PCIe both ways every step; per-cell rate on the first cells of the workload"}
        chk = min(rt_cells, 4096)
        S_dev = moc.source(n0[:, :chk].contiguous(), eps[:chk].contiguous(), rho[:chk].contiguous(),
                           nu[:chk].contiguous(), w.dt)
        if not torch.equal(S_dev.cpu(), S_host[:, :chk]):
            raise SystemExit("bench.py: host-buffer path and resident path disagree")
    del n0, n_host, S_host
    out = {"value": value, "ms_per_step": ms_per_step, "warmup": warmup, "launches": int(launches), "clocks": clocks,
           "roofline": roofline, "e2e": e2e, "strong": strong, "w": w, "cells": cells, "moc": moc,
           "checks": {"n_min_after_steps": n_min, "global_moments": gmom, "input_fill_s": t_gen}}
    return out


def leg_config1(B: Bench):
    """BASELINE configs[0]: MOC 30 classes, 1e5 cells, benchmark constants (device + resident e2e)."""
    torch = B.torch
    from openpbe_b200 import engine, synthetic

    w = synthetic.workload("moc30_ct")
    cells = w.ncells
    f = synthetic.moc_fields(w, offset=B.rank * cells)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(B.ctx, w.N, w.xi1, spec, w.relax)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(B.dev)  # noqa: E731
    n0, eps, rho, nu = T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"])
    n = n0.clone()
    for _ in range(3):
        n.copy_(n0)
        moc.step(n, eps, rho, nu, w.dt)
    B.barrier()
    tot, steps = 0.0, max(5, B.args.steps)
    for _ in range(steps):
        n.copy_(n0)
        B.flush_l2()  # 50 MB of fields would otherwise sit in the 126 MB L2
        B.e0.record()
        moc.step(n, eps, rho, nu, w.dt)
        B.e1.record()
        torch.cuda.synchronize()
        tot += B.e0.elapsed_time(B.e1)
    ms = B.allmax(tot / steps)
    flops, byts = moc_flops_per_cell(w), moc_bytes_per_cell(w.N)
    tf = cells * flops / (ms * 1e-3) / 1e12
    gbs = cells * byts / (ms * 1e-3) / 1e9
    t_fp, t_hbm = cells * flops / (B.fp64_tf * 1e12), cells * byts / (B.hbm_peak * 1e9)
    kinfo = moc.kernel()
    waves = cells / (kinfo["cells_per_cta"] * kinfo["ctas_per_sm"] * B.info["sm_count"])
    leg = {"metric": "moc30_cell_class_updates_per_s", "value": B.world * cells * w.N / (ms * 1e-3), "unit": UNIT,
           "ms_per_step": ms, "cells_per_gpu": cells, "classes": w.N,
           "workload": "configs[0]: MOC 30 classes x 1e5 cells, CT coalescence + CT breakup with the constants of "
                       "benchmarks/coalescence.cpp:28-33 / breakup.cpp:29-33, one in-place zero-flux step",
           "l2": "flushed between timed launches (256 MB memset)",
           "roofline": {"bound": "fp64", "achieved": tf, "peak": B.fp64_tf, "unit": "TFLOP/s", "frac": tf / B.fp64_tf,
                        "t_star_ms": max(t_fp, t_hbm) * 1e3, "frac_of_t_star": max(t_fp, t_hbm) * 1e3 / ms,
                        "hbm_gbs": gbs, "flops_per_cell": flops, "traffic": None,
                        "kernel": f"{kinfo['kernel']} ({kinfo['cells_per_cta']} cells/CTA, {kinfo['ctas_per_sm']} CTA/SM)",
                        "note": f"1e5 cells = {waves:.2f} waves of the persistent grid: the launch is a single "
                                "partial wave, so the tail and the launch latency are inside the time"}}
    if not B.args.no_e2e:
        st = engine.MOCState(moc, cells, vcell=f["vcell"], n0=f["n"])
        hp = {k: B.pinned(f[k]) for k in ("eps", "rho", "nu", "alpha")}
        dh = torch.empty(cells, dtype=torch.float64, pin_memory=True)
        for _ in range(2):
            st.correct(hp["eps"], hp["rho"], hp["nu"], hp["alpha"], w.dt, d_out=dh, minmax=True)
        t0 = time.perf_counter()
        for _ in range(10):
            st.correct(hp["eps"], hp["rho"], hp["nu"], hp["alpha"], w.dt, d_out=dh, minmax=True)
        te = B.allmax((time.perf_counter() - t0) / 10)
        leg["e2e"] = {"value": B.world * cells * w.N / te, "unit": UNIT, "ms_per_step": te * 1e3,
                      "h2d_bytes_per_step": 32 * cells, "d2h_bytes_per_step": 8 * cells,
                      "call": "pbe_moc_correct_resident (pinned host eps, rho, nu, alpha -> d, per-class min)"}
        st.close()
    moc.close()
    return leg, w


def leg_qmom(B: Bench):
    """The second half of the metric: QMOM 6 moments, BASELINE configs[1]."""
    torch, args = B.torch, B.args
    from openpbe_b200 import engine, synthetic

    wq = synthetic.workload("qmom3_ct")
    qc = int(args.qmom_cells) if args.qmom_cells else wq.ncells
    mh = torch.empty((2 * wq.N, qc), dtype=torch.float64)
    fq_small = [torch.empty(qc, dtype=torch.float64) for _ in range(3)]
    blk = 1_000_000
    for b0 in range(0, qc, blk):
        nb = min(blk, qc - b0)
        fq = synthetic.qmom_fields(wq, offset=B.rank * qc + b0, ncells=nb)
        mh.numpy()[:, b0:b0 + nb] = fq["m"]
        for t, k in zip(fq_small, ("eps", "rho", "nu")):
            t.numpy()[b0:b0 + nb] = fq[k]
    m0 = mh.to(B.dev)
    qe, qr, qn = (t.to(B.dev) for t in fq_small)
    qspec = engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs)
    q = engine.QMOM(B.ctx, wq.N, qspec)
    dq = torch.empty(qc, dtype=torch.float64, device=B.dev)
    mq = m0.clone()

    # what QMOM::correct() needs per step (QMOM.C:85-129): m updated in place and d; the scrubbed
    # sources are only printed (printAvgMaxMin), so the timed step does not store them
    def qstep():
        q.step(mq, qe, qr, qn, wq.dt, d=dq)

    # every step starts from the synthetic state (restored outside the timed region): with
    # quadrature weights of 1e5..1e7 the explicit update leaves most moments on the SMALL clamp
    # after one step (QMOM.C:117-123), and later steps would time the degenerate-cell path
    for _ in range(max(3, args.warmup)):
        mq.copy_(m0)
        qstep()
    B.barrier()
    tot = 0.0
    qsteps = max(3, args.steps)
    for _ in range(qsteps):
        mq.copy_(m0)
        B.flush_l2()
        B.e0.record()
        qstep()
        B.e1.record()
        torch.cuda.synchronize()
        tot += B.e0.elapsed_time(B.e1)
    qms = B.allmax(tot / qsteps)
    qbytes = 8.0 * (6 + 6 + 3 + 1)  # read 6 m + eps,rho,nu; write 6 m, d
    qgbs = qc * qbytes / (qms * 1e-3) / 1e9
    qtf = qc * FLOPS_QMOM3_CT / (qms * 1e-3) / 1e12
    t_hbm, t_fp = qc * qbytes / (B.hbm_peak * 1e9), qc * FLOPS_QMOM3_CT / (B.fp64_tf * 1e12)
    bound = "fp64" if t_fp >= t_hbm else "hbm"
    traffic, tsrc = load_traffic("qmom_kernel", qc)
    leg = {"metric": "qmom6_cell_updates_per_s", "value": B.world * qc / (qms * 1e-3), "unit": "cell-updates/s",
           "ms_per_step": qms, "cells_per_gpu": qc,
           "workload": "qmom3_ct: Wheeler inversion + 6 moment sources + scrub/update/clamp/d, CT kernels (fused "
                       "pbe_qmom_step, m in place + d out)",
           "l2": "flushed between timed launches (256 MB memset)",
           "roofline": {"bound": bound, "t_star_ms": max(t_hbm, t_fp) * 1e3, "frac": max(t_hbm, t_fp) * 1e3 / qms,
                        "achieved": qtf if bound == "fp64" else qgbs,
                        "peak": B.fp64_tf if bound == "fp64" else B.hbm_peak,
                        "unit": "TFLOP/s" if bound == "fp64" else "GB/s",
                        "hbm": {"achieved": qgbs, "peak": B.hbm_peak, "unit": "GB/s", "frac": qgbs / B.hbm_peak,
                                "bytes_per_cell": qbytes},
                        "fp64": {"achieved": qtf, "peak": B.fp64_tf, "unit": "TFLOP/s", "frac": qtf / B.fp64_tf,
                                 "flops_per_cell": FLOPS_QMOM3_CT},
                        "traffic": traffic, "traffic_source": f"committed constant ({tsrc}), not measured by this run",
                        "algorithmic_bytes": qc * qbytes,
                        "note": "T* = max(bytes / HBM peak, flops / FP64 peak), SURVEY.md 8(d)"}}
    if not args.no_e2e:
        st = engine.QMOMState(q, qc, m0=mh)
        sp = [t.pin_memory() for t in fq_small]
        dh = torch.empty(qc, dtype=torch.float64, pin_memory=True)
        sm, _ = st.device_fields()
        for _ in range(2):
            sm.copy_(m0)
            st.correct(*sp, wq.dt, d_out=dh)
        B.barrier()
        tq = 0.0
        for _ in range(3):
            sm.copy_(m0)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            st.correct(*sp, wq.dt, d_out=dh)
            tq += time.perf_counter() - t0
        tq = B.allmax(tq / 3)
        leg["e2e"] = {"value": B.world * qc / tq, "unit": "cell-updates/s", "ms_per_step": tq * 1e3,
                      "h2d_bytes_per_step": 24 * qc, "d2h_bytes_per_step": 8 * qc, "pcie_gbs": 32 * qc / tq / 1e9,
                      "call": "pbe_qmom_correct_resident: moments owned on the device (QMOM.H:62); pinned host eps, "
                              "rho, nu -> d; synchronous"}
        del sm
        st.close()
    q.close()
    return leg, wq


def leg_replay(B: Bench):
    """BASELINE configs[4]: PBEMethod::correct() per time step on synthetic 4e6-cell fields, MOC
    (50 classes, c01 constants) and QMOM (order 3) side by side, through the resident correct()."""
    torch, args = B.torch, B.args
    from openpbe_b200 import engine, synthetic

    out = {"workload": "configs[4]: correct() replay with the constants of tests/validation/simmons_and_azzopardi/c01 "
                       "(phaseProperties:24-50), zero flux, 4e6 synthetic cells per GPU; per step the pinned host "
                       "phase fields go in and d comes back, as PBEFoam.C:56-67 drives it",
           "steps": int(args.replay_steps)}
    wm = synthetic.workload("replay_moc")
    cells = int(args.replay_cells)
    spec = engine.make_spec(wm.coalescence, wm.coal_coeffs, wm.breakup, wm.br_coeffs)
    moc = engine.MOC(B.ctx, wm.N, wm.xi1, spec, wm.relax)
    hp = {k: torch.empty(cells, dtype=torch.float64, pin_memory=True) for k in ("eps", "rho", "nu", "alpha", "vcell")}
    n_tmp = torch.empty((wm.N, cells), dtype=torch.float64, device=B.dev)
    blk = 500_000
    for c0 in range(0, cells, blk):
        nb = min(blk, cells - c0)
        f = synthetic.moc_fields(wm, offset=B.rank * cells + c0, ncells=nb)
        n_tmp[:, c0:c0 + nb].copy_(torch.from_numpy(f["n"]))
        for k in hp:
            hp[k][c0:c0 + nb].copy_(torch.from_numpy(f[k]))
    st = engine.MOCState(moc, cells, vcell=hp["vcell"])
    sn, _ = st.device_fields()
    sn.copy_(n_tmp)
    torch.cuda.synchronize()
    del n_tmp
    dh = torch.empty(cells, dtype=torch.float64, pin_memory=True)
    for _ in range(2):
        st.correct(hp["eps"], hp["rho"], hp["nu"], hp["alpha"], wm.dt, d_out=dh, minmax=True)
    B.barrier()
    l0 = B.ctx.launch_count()
    t0 = time.perf_counter()
    for _ in range(int(args.replay_steps)):
        d, diag = st.correct(hp["eps"], hp["rho"], hp["nu"], hp["alpha"], wm.dt, d_out=dh, minmax=True)
    tm = B.allmax((time.perf_counter() - t0) / int(args.replay_steps))
    nl = (B.ctx.launch_count() - l0) // int(args.replay_steps)
    # kernel-only time of the same step for the roofline
    eps_d, rho_d, nu_d = (hp[k].to(B.dev) for k in ("eps", "rho", "nu"))
    B.e0.record()
    for _ in range(3):
        moc.step(sn, eps_d, rho_d, nu_d, wm.dt)
    B.e1.record()
    torch.cuda.synchronize()
    kms = B.allmax(B.e0.elapsed_time(B.e1) / 3)
    flops = moc_flops_per_cell(wm)
    out["moc"] = {"method": "methodOfClasses", "classes": wm.N, "cells_per_gpu": cells, "ms_per_step": tm * 1e3,
                  "value": B.world * cells * wm.N / tm, "unit": UNIT, "launches_per_step": int(nl),
                  "h2d_bytes_per_step": 32 * cells, "d2h_bytes_per_step": 8 * cells,
                  "d_mean": float(d.mean()), "min_n_last_class": float(diag["class_min"][-1]),
                  "cont_err_max": diag["cont_err_max"], "kernel_ms_per_step": kms,
                  "roofline": {"bound": "fp64", "achieved": cells * flops / (kms * 1e-3) / 1e12, "peak": B.fp64_tf,
                               "unit": "TFLOP/s", "frac": cells * flops / (kms * 1e-3) / 1e12 / B.fp64_tf,
                               "flops_per_cell": flops, "traffic": None}}
    del sn, eps_d, rho_d, nu_d
    st.close()
    moc.close()

    wq = synthetic.workload("replay_qmom")
    q = engine.QMOM(B.ctx, wq.N, engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs))
    m0 = torch.empty((2 * wq.N, cells), dtype=torch.float64)
    qh = {k: torch.empty(cells, dtype=torch.float64, pin_memory=True) for k in ("eps", "rho", "nu")}
    for c0 in range(0, cells, 1_000_000):
        nb = min(1_000_000, cells - c0)
        f = synthetic.qmom_fields(wq, offset=B.rank * cells + c0, ncells=nb)
        m0[:, c0:c0 + nb] = torch.from_numpy(f["m"])
        for k in qh:
            qh[k][c0:c0 + nb] = torch.from_numpy(f[k])
    qs = engine.QMOMState(q, cells, m0=m0)
    for _ in range(2):
        qs.correct(qh["eps"], qh["rho"], qh["nu"], wq.dt, d_out=dh)
    B.barrier()
    t0 = time.perf_counter()
    for _ in range(int(args.replay_steps)):
        dq, _ = qs.correct(qh["eps"], qh["rho"], qh["nu"], wq.dt, d_out=dh)
    tq = B.allmax((time.perf_counter() - t0) / int(args.replay_steps))
    out["qmom"] = {"method": "quadratureMethodOfMoments", "order": wq.N, "cells_per_gpu": cells,
                   "ms_per_step": tq * 1e3, "value": B.world * cells / tq, "unit": "cell-updates/s",
                   "h2d_bytes_per_step": 24 * cells, "d2h_bytes_per_step": 8 * cells, "pcie_gbs": 32 * cells / tq / 1e9,
                   "d_mean": float(dq.mean()),
                   "note": "PCIe-bound: 32 B/cell over the bus against 0.03 ms of kernel per 1e6 cells"}
    qs.close()
    q.close()
    return out, wm, wq


def leg_mom(B: Bench):
    torch, args = B.torch, B.args
    from openpbe_b200 import engine

    mc = int(args.mom_cells)
    rng = np.random.default_rng(20240611 + B.rank)
    gm0, gk, gth = rng.uniform(0.5, 3.0, mc), rng.uniform(1.5, 4.0, mc), rng.uniform(0.15, 0.6, mc)
    mm0 = torch.from_numpy(np.stack([gm0, gm0 * gk * gth, gm0 * gk * (gk + 1) * gth ** 2])).to(B.dev)
    me, mr, mn = (torch.from_numpy(rng.uniform(lo, hi, mc)).to(B.dev)
                  for lo, hi in ((0.5, 2.0), (0.8, 1.2), (0.5, 1.5)))
    ct1 = (1.0, 1.0, 0.0, 1.0)  # O(1) scales: where integrate()'s epsabs 1e-7 means something
    mom = engine.MOM(B.ctx, engine.make_spec("CoulaloglouTavlaridesC", ct1, "CoulaloglouTavlarides", ct1))
    mwork, md = mm0.clone(), torch.empty(mc, dtype=torch.float64, device=B.dev)
    for _ in range(3):
        mwork.copy_(mm0)
        mom.step(mwork, me, mr, mn, 1e-3, d=md)
    B.barrier()
    tot, msteps = 0.0, min(5, max(3, args.steps))
    for _ in range(msteps):
        mwork.copy_(mm0)
        B.e0.record()
        mom.step(mwork, me, mr, mn, 1e-3, d=md)
        B.e1.record()
        torch.cuda.synchronize()
        tot += B.e0.elapsed_time(B.e1)
    mms = B.allmax(tot / msteps)
    nq = 15 * (12 + 20)
    # convention of SURVEY 8(d): exp = 32, sqrt / div = 1, FMA = 2; per unordered node pair of the CT
    # kernel 4 add + 7 mul + sqrt + div + exp + 2 FMA = 49; per node the gamma pdf (log, exp, 4) = 68
    mflops = 49.0 * (nq * (nq + 1) / 2) + 68.0 * nq
    mtf = mc * mflops / (mms * 1e-3) / 1e12
    return {"metric": "mom3_cell_updates_per_s", "value": B.world * mc / (mms * 1e-3), "unit": "cell-updates/s",
            "ms_per_step": mms, "cells_per_gpu": mc, "quadrature_nodes": nq,
            "workload": "methodOfMoments: gamma closure + coalescence / breakup source integrals (CT kernels, "
                        "O(1) scales) on the fixed graded Gauss-Kronrod rule + scrub/update/clamp/d",
            "l2": "compute-bound: 56 B per cell of HBM traffic, no flush needed",
            "roofline": {"bound": "fp64", "achieved": mtf, "peak": B.fp64_tf, "unit": "TFLOP/s",
                         "frac": mtf / B.fp64_tf, "flops_per_cell": mflops, "traffic": None,
                         "note": "flop convention of SURVEY 8(d) applied to the reference's arithmetic per node pair"}}


def leg_geometric(B: Bench, w):
    """BASELINE configs[2] as worded ("100 geometric classes"): the fixed-pivot extension (row f4, parity
    unpinned) on a geometric grid, beside the headline which runs the reference's uniform grid."""
    torch, args = B.torch, B.args
    from openpbe_b200 import engine, synthetic

    gc, ratio = int(args.geo_cells), float(args.geo_ratio)
    x = w.xi1 * ratio ** np.arange(w.N)
    f = synthetic.moc_fields(w, ncells=min(gc, 200_000))
    rep = (gc + f["n"].shape[1] - 1) // f["n"].shape[1]
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(B.dev)  # noqa: E731
    n0 = T(np.tile(f["n"], (1, rep))[:, :gc])
    n = n0.clone()
    e, r, v = (T(np.tile(f[k], rep)[:gc]) for k in ("eps", "rho", "nu"))
    moc = engine.MOC(B.ctx, w.N, 0.0, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs),
                     w.relax, pivots=x)
    mom = torch.zeros(6, dtype=torch.float64, device=B.dev)
    for _ in range(2):
        moc.step(n, e, r, v, w.dt, moments=mom)
    B.barrier()
    gsteps = 3
    B.e0.record()
    for _ in range(gsteps):
        moc.step(n, e, r, v, w.dt, moments=mom)
    B.e1.record()
    torch.cuda.synchronize()
    gms = B.allmax(B.e0.elapsed_time(B.e1) / gsteps)
    top = x[-1] * (1 + 2e-15)
    npairs = int(sum(np.searchsorted(x, top - x[i], side="right") - i for i in range(w.N) if 2 * x[i] <= top))
    gflops = 40.0 + 43.0 * npairs + 40.0 * w.N  # SURVEY 8(d) convention + the two-pivot split (mul, sub) per pair
    gtf = gc * gflops / (gms * 1e-3) / 1e12
    info = moc.kernel()
    ge2e = None
    if not args.no_e2e:
        # the same step end to end: class fields owned on the device, pinned host eps, rho, nu, alpha -> d
        st = engine.MOCState(moc, gc, vcell=np.tile(f["vcell"], rep)[:gc])
        sn, _ = st.device_fields()
        sn.copy_(n0)
        hb = {k: B.pinned(np.tile(f[k], rep)[:gc]) for k in ("eps", "rho", "nu", "alpha")}
        dh = torch.empty(gc, dtype=torch.float64, pin_memory=True)
        st.correct(hb["eps"], hb["rho"], hb["nu"], hb["alpha"], w.dt, d_out=dh, minmax=True, n_moments=6)
        t0 = time.perf_counter()
        for _ in range(3):
            st.correct(hb["eps"], hb["rho"], hb["nu"], hb["alpha"], w.dt, d_out=dh, minmax=True, n_moments=6)
        te = (time.perf_counter() - t0) / 3
        ge2e = {"value": gc * w.N / te, "unit": "cell-class-updates/s", "ms_per_step": te * 1e3,
                "h2d_bytes_per_step": 32 * gc, "d2h_bytes_per_step": 8 * gc,
                "call": "pbe_moc_correct_resident on a fixed-pivot handle (zero flux)"}
        del sn
        st.close()
    moc.close()
    return {"metric": "moc_cell_class_updates_per_s", "value": B.world * gc * w.N / (gms * 1e-3), "e2e": ge2e,
            "unit": "cell-class-updates/s", "ms_per_step": gms, "cells_per_gpu": gc, "classes": w.N,
            "grid": f"geometric, ratio {ratio}: class volumes xi1 {ratio}^i", "pairs_per_cell": npairs,
            "workload": "MOC fixed-pivot step (Kumar-Ramkrishna two-pivot redistribution in the reference's "
                        "conventions; extension, parity unpinned), CT + CT kernels, c01 constants, in-place step + "
                        "6 fused moments", "kernel": info,
            "roofline": {"bound": "fp64", "achieved": gtf, "peak": B.fp64_tf, "unit": "TFLOP/s",
                         "frac": gtf / B.fp64_tf, "flops_per_cell": gflops, "traffic": None,
                         "note": "4x4 pair tiles through a TMA-fed ring, register birth window, accumulators in tensor memory: two CTAs per SM (DESIGN.md 4.8)"}}


def leg_transport(B: Bench, w, moc):
    torch, args = B.torch, B.args
    from openpbe_b200 import capi, engine, sharding, synthetic

    rank, world, dev, ctx = B.rank, B.world, B.dev, B.ctx
    N = w.N
    tn = int(args.transport_n)
    # one n^3 block per GPU: z-slabs of an n x n x (n world) block, processor patches between them
    tm = synthetic.block_slab_part(tn, rank, world, velocity=(1.0, 0.5, 0.25))
    tmesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    tmesh.set_coupled(*tm["coupled"])
    halo_x, halo_kind = None, None
    if world > 1:
        halo_nccl = sharding.HaloExchange(tm, dev, mesh=tmesh)
        halo_x, halo_kind = halo_nccl, "NCCL send/recv"
        if not args.halo_nccl:
            try:  # pack kernel storing straight into the neighbour's symmetric-memory buffer
                halo_x = sharding.PeerHaloExchange(tm, dev, tmesh, N)
                halo_kind = "pack kernel stores into peer (symmetric) memory + one device barrier"
            except Exception as ex:  # noqa: BLE001 -- reported, not hidden
                halo_kind = f"NCCL send/recv (peer-memory exchange unavailable: {type(ex).__name__}: {ex})"[:300]
    tinfo = tmesh.info()
    tc, tK = tn ** 3, N
    gen = torch.Generator(device=dev)
    gen.manual_seed(20240611 + rank)
    tpsi = torch.rand(tK, tc, dtype=torch.float64, device=dev, generator=gen)
    tS = torch.rand(tK, tc, dtype=torch.float64, device=dev, generator=gen)
    tout = torch.empty_like(tpsi)
    tphi = torch.from_numpy(tm["phi"]).to(dev)
    tdt = 0.4 / (tn * 1.75)  # CFL 0.4

    def tstep():
        if halo_x is None:
            tmesh.transport_explicit(tpsi, tphi, tdt, S=tS, out=tout)
            return
        # interior (regular) cells read no halo value: their update overlaps the exchange
        h = halo_x.start(tpsi)
        tmesh.transport_explicit(tpsi, tphi, tdt, S=tS, out=tout, cells=capi.CELLS_REGULAR)
        halo = halo_x.finish(h)
        tmesh.transport_explicit(tpsi, tphi, tdt, S=tS, out=tout, psi_halo=halo, cells=capi.CELLS_IRREGULAR)

    if halo_x is not None and halo_x is not halo_nccl:
        # the peer-memory exchange must deliver exactly what the NCCL exchange delivers
        for _ in range(3):  # both buffer parities
            a = halo_x.exchange(tpsi).clone()
            b = halo_nccl.exchange(tpsi)
            if not torch.equal(a, b):
                raise SystemExit("bench.py: peer-memory halo exchange differs from the NCCL exchange")
    l0 = ctx.launch_count()
    for _ in range(max(3, args.warmup)):
        tstep()
    t_launches = (ctx.launch_count() - l0) // max(3, args.warmup)
    B.barrier()
    tsteps = max(3, args.steps)
    B.e0.record()
    for _ in range(tsteps):
        tstep()
    B.e1.record()
    torch.cuda.synchronize()
    tms = B.allmax(B.e0.elapsed_time(B.e1) / tsteps)
    # psi and S in, psi out per cell-field; V, 6 x (neighbour, face) labels and 3 face fluxes per cell
    tbytes = tc * (24.0 * tK + 8 + 6 * 8 + 3 * 8)
    tgbs = tbytes / (tms * 1e-3) / 1e9
    traffic, tsrc = load_traffic("transport_upwind_kernel", tc)
    transport = {"metric": "transport_cell_field_updates_per_s", "value": world * tc * tK / (tms * 1e-3),
                 "unit": "cell-field-updates/s", "ms_per_step": tms, "cells_per_gpu": tc, "fields": tK,
                 "workload": f"explicit upwind transport + source update (MULES::explicitSolve form) of {tK} "
                             f"class fields on a {tn}^3 hex block in OpenFOAM face addressing, uniform "
                             "velocity (1, 0.5, 0.25), CFL 0.4; one block per GPU"
                             + (f", z-slabs of one {tn}x{tn}x{tn * world} block: halo exchange of "
                                f"{tK} x {tn * tn} values per processor patch ({halo_kind}) inside the "
                                "timed step, overlapped with the interior cells' update" if world > 1 else ""),
                 "l2": f"fields ({3 * tc * tK * 8 / 1e9:.1f} GB) exceed the 126 MB L2",
                 "mesh": tinfo, "launches_per_step": int(t_launches),
                 "roofline": {"bound": "hbm", "achieved": tgbs, "peak": B.hbm_peak, "unit": "GB/s",
                              "frac": tgbs / B.hbm_peak, "algorithmic_bytes": tbytes, "traffic": traffic,
                              "traffic_source": f"committed constant ({tsrc}), not measured by this run"}}
    if world == 1 and "Sf" in tm:
        # the limited path of MOC::solveWithMULES (MOC.C:198-272): Gauss vanLeer face values, MULES limiter
        # (3 sweeps), limitSum over the fields, explicitSolve -- the same fields, mesh and time step
        tmesh.set_geometry(tm["Sf"], tm["C"], tm["weights"])
        pmax = torch.full((tK,), 2.0, dtype=torch.float64, device=dev)
        for _ in range(2):
            tmesh.transport_mules(tpsi, tphi, tdt, pmax, S=tS, out=tout)
        l0 = ctx.launch_count()
        B.e0.record()
        for _ in range(3):
            tmesh.transport_mules(tpsi, tphi, tdt, pmax, S=tS, out=tout)
        B.e1.record()
        torch.cuda.synchronize()
        mms = B.e0.elapsed_time(B.e1) / 3
        transport["mules"] = {
            "metric": "transport_cell_field_updates_per_s", "value": tc * tK / (mms * 1e-3),
            "unit": "cell-field-updates/s", "ms_per_step": mms, "scheme": "Gauss vanLeer, nLimiterIter 3",
            "launches_per_step": int((ctx.launch_count() - l0) // 3),
            "roofline": {"bound": "hbm", "achieved": tbytes / (mms * 1e-3) / 1e9, "peak": B.hbm_peak, "unit": "GB/s",
                         "frac": tbytes / (mms * 1e-3) / 1e9 / B.hbm_peak, "algorithmic_bytes": tbytes,
                         "note": "same algorithmic bytes as the upwind update (psi and S in, psi out): the gradient, "
                                 "correction-flux, limiter-sweep and limitSum passes in between are what the "
                                 "fraction pays for", "traffic": None}}
        del pmax
    if world == 1:
        # MOC::solveWithFVM (MOC.C:186-198), the reference's DEFAULT branch (usingMULES false): fvm::ddt + fvm::div == S
        # of all fields at once, stopped by the fvSolution of the reference's cases (tolerance 1e-5, relTol 0, minIter 1)
        for _ in range(2):
            _, perf = tmesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout)
        l0 = ctx.launch_count()
        B.e0.record()
        for _ in range(3):
            tmesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout)
        B.e1.record()
        torch.cuda.synchronize()
        ims = B.e0.elapsed_time(B.e1) / 3
        its = [p[2] for p in perf]
        # set-up: psi and S in, x and z out (32 B per cell-field); every Jacobi sweep of a field: z in, x in / out,
        # z out (32 B per cell-field); the per-cell rows are small beside them
        ibytes = tc * 32.0 * (tK + sum(its)) + tc * (8 + 6 * 8 + 3 * 8 + 4 * 8)
        transport["implicit"] = {
            "metric": "transport_cell_field_updates_per_s", "value": tc * tK / (ims * 1e-3),
            "unit": "cell-field-updates/s", "ms_per_step": ims, "scheme": "Euler implicit + Gauss upwind",
            "controls": {"tolerance": 1e-5, "relTol": 0, "minIter": 1, "solver": "Jacobi (auto)"},
            "iterations": [min(its), max(its)], "initial_residual_max": max(p[0] for p in perf),
            "final_residual_max": max(p[1] for p in perf),
            "launches_per_step": int((ctx.launch_count() - l0) // 3),
            "roofline": {"bound": "hbm", "achieved": ibytes / (ims * 1e-3) / 1e9, "peak": B.hbm_peak, "unit": "GB/s",
                         "frac": ibytes / (ims * 1e-3) / 1e9 / B.hbm_peak, "algorithmic_bytes": ibytes,
                         "note": "bytes of the iteration actually run: 32 B per cell-field for the set-up and for "
                                 "every sweep a field needed (DESIGN.md 4.11)", "traffic": None}}
        if "Sf" in tm:  # the scheme of the reference's cases in the implicit term: one matrix per field
            for _ in range(2):
                _, perf = tmesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout, scheme="Gauss vanLeer")
            B.e0.record()
            for _ in range(3):
                tmesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout, scheme="Gauss vanLeer")
            B.e1.record()
            torch.cuda.synchronize()
            vms = B.e0.elapsed_time(B.e1) / 3
            transport["implicit"]["vanLeer"] = {"ms_per_step": vms, "value": tc * tK / (vms * 1e-3),
                                                "unit": "cell-field-updates/s", "scheme": "Euler implicit + Gauss vanLeer",
                                                "iterations": [min(p[2] for p in perf), max(p[2] for p in perf)],
                                                "final_residual_max": max(p[1] for p in perf)}
    if world > 1:
        # the same implicit solve on the decomposed block: the ranks solve ONE system -- z packed and exchanged per
        # Jacobi sweep (ncclSend / ncclRecv), the per-field sums all-reduced, all through the C ABI's communicator
        # (pbe_comm_init) on a context of its own
        import ctypes as C
        import torch.distributed as dist
        ident = C.create_string_buffer(128)
        if rank == 0:
            capi.check(capi.lib().pbe_comm_unique_id(ident))
        box = [bytes(ident.raw)]
        dist.broadcast_object_list(box, src=0)
        ident = C.create_string_buffer(box[0], 128)
        ictx = engine.Context(dev.index)
        capi.check(capi.lib().pbe_comm_init(ictx.handle, world, rank, ident))
        imesh = engine.Mesh(ictx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
        imesh.set_coupled(*tm["coupled"])
        imesh.set_halo_plan(tm["neighbours"], tm["send_cells"], tc * world)
        for _ in range(2):
            _, perf = imesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout)
        l0 = ictx.launch_count()
        B.barrier()
        B.e0.record()
        for _ in range(3):
            imesh.transport_implicit(tpsi, tphi, tdt, S=tS, out=tout)
        B.e1.record()
        torch.cuda.synchronize()
        ims = B.allmax(B.e0.elapsed_time(B.e1) / 3)
        its = [p[2] for p in perf]
        transport["implicit"] = {
            "metric": "transport_cell_field_updates_per_s", "value": world * tc * tK / (ims * 1e-3),
            "unit": "cell-field-updates/s", "ms_per_step": ims, "scheme": "Euler implicit + Gauss upwind",
            "controls": {"tolerance": 1e-5, "relTol": 0, "minIter": 1, "solver": "Jacobi (auto)"},
            "iterations": [min(its), max(its)], "final_residual_max": max(p[1] for p in perf),
            "launches_per_step": int((ictx.launch_count() - l0) // 3), "ranks": world,
            "roofline": (lambda by: {"bound": "hbm", "achieved": by / (ims * 1e-3) / 1e9, "peak": B.hbm_peak, "unit": "GB/s",
                                     "frac": by / (ims * 1e-3) / 1e9 / B.hbm_peak, "algorithmic_bytes": by, "traffic": None,
                                     "note": "per GPU: 32 B per cell-field for the set-up and for every sweep a field "
                                             "needed; the halo exchange and the all-reduce are inside the time"})(
                tc * 32.0 * (tK + sum(its))),
            "decomposition": f"z-slabs of one {tn}x{tn}x{tn * world} block; per sweep {tK} x {tn * tn} values per processor "
                             "patch packed and exchanged by ncclSend / ncclRecv, 3 x nfields sums all-reduced, on the "
                             "solve's stream (pbe_mesh_set_halo_plan)"}
        imesh.close()
        capi.check(capi.lib().pbe_comm_destroy(ictx.handle))
        ictx.close()
    if not args.no_e2e and world == 1:
        # MOC::correct() with solveWithMULES through the resident state: class fields stay on the device,
        # pinned host eps, rho, nu, alpha, phi in, d out
        ft = synthetic.moc_fields(w, ncells=min(tc, 200_000))
        rep = (tc + ft["n"].shape[1] - 1) // ft["n"].shape[1]
        st = engine.MOCState(moc, tc, vcell=tm["vcell"], mesh=tmesh)
        sn, _ = st.device_fields()
        sn.copy_(torch.from_numpy(np.tile(ft["n"], (1, rep))[:, :tc].copy()))
        sm = {k: B.pinned(np.tile(ft[k], rep)[:tc]) for k in ("eps", "rho", "nu", "alpha")}
        ph = B.pinned(tm["phi"])
        dh = torch.empty(tc, dtype=torch.float64, pin_memory=True)
        st.correct(sm["eps"], sm["rho"], sm["nu"], sm["alpha"], w.dt, phi=ph, d_out=dh, minmax=True)
        t0 = time.perf_counter()
        for _ in range(3):
            st.correct(sm["eps"], sm["rho"], sm["nu"], sm["alpha"], w.dt, phi=ph, d_out=dh, minmax=True)
        te = (time.perf_counter() - t0) / 3
        transport["e2e"] = {"value": tc * tK / te, "unit": "cell-class-updates/s", "ms_per_step": te * 1e3,
                            "h2d_bytes_per_step": 32 * tc + tm["phi"].size * 8, "d2h_bytes_per_step": 8 * tc,
                            "call": "pbe_moc_correct_resident with an attached mesh (pinned host eps, rho, nu, alpha, "
                                    "phi -> d): sources + upwind transport update of the resident class fields",
                            "d_mean": float(dh.mean())}
        # the same call with usingMULES false, the reference's default: sources + solveWithFVM (implicit, fvSolution of the
        # reference's cases); deltaT is the transport legs' (Courant 0.4) so that the solve is the one timed above
        st.set_implicit(tolerance=1e-5, min_iter=1)
        st.correct(sm["eps"], sm["rho"], sm["nu"], sm["alpha"], tdt, phi=ph, d_out=dh, minmax=True)
        t0 = time.perf_counter()
        for _ in range(3):
            st.correct(sm["eps"], sm["rho"], sm["nu"], sm["alpha"], tdt, phi=ph, d_out=dh, minmax=True)
        ti = (time.perf_counter() - t0) / 3
        pf = st.solver_performance()
        transport["e2e_implicit"] = {"value": tc * tK / ti, "unit": "cell-class-updates/s", "ms_per_step": ti * 1e3,
                                     "h2d_bytes_per_step": 32 * tc + tm["phi"].size * 8, "d2h_bytes_per_step": 8 * tc,
                                     "iterations": [min(p[2] for p in pf), max(p[2] for p in pf)],
                                     "call": "pbe_moc_correct_resident after pbe_moc_state_set_implicit: sources + the "
                                             "implicit solve of all class fields (MOC::solveWithFVM)",
                                     "d_mean": float(dh.mean())}
        del sn
        st.close()
    del tpsi, tS, tout, tphi, tmesh
    return transport


def run_b200(args):
    B = Bench(args)
    from openpbe_b200 import synthetic

    rank, world = B.rank, B.world
    main = leg_main(B)
    w, cells, N = main["w"], main["cells"], main["w"].N
    roofline, e2e = main["roofline"], main["e2e"]
    config1 = qmom = replay = transport = mom_leg = None
    w1 = wq = wrm = wrq = None
    if not args.no_config1:
        config1, w1 = leg_config1(B)
        roofline["legs"]["config1_moc30"] = config1["roofline"]
        if e2e is not None and "e2e" in config1:
            e2e["legs"]["config1_moc30"] = config1["e2e"]
    if not args.no_qmom:
        qmom, wq = leg_qmom(B)
        roofline["legs"]["qmom"] = qmom["roofline"]
        if e2e is not None and "e2e" in qmom:
            e2e["legs"]["qmom"] = qmom["e2e"]
    if not args.no_replay and not args.no_e2e:
        replay, wrm, wrq = leg_replay(B)
        roofline["legs"]["replay_moc50"] = replay["moc"]["roofline"]
        e2e["legs"]["replay_moc50"] = {k: replay["moc"][k] for k in ("ms_per_step", "value", "unit",
                                                                      "h2d_bytes_per_step", "d2h_bytes_per_step")}
        e2e["legs"]["replay_qmom3"] = {k: replay["qmom"][k] for k in ("ms_per_step", "value", "unit",
                                                                       "h2d_bytes_per_step", "d2h_bytes_per_step")}
    if not args.no_mom:
        mom_leg = leg_mom(B)
        roofline["legs"]["mom"] = mom_leg["roofline"]
    geometric = None
    if not args.no_geometric:
        geometric = leg_geometric(B, w)
        roofline["legs"]["moc100_geometric"] = geometric["roofline"]
        if e2e is not None and geometric.get("e2e"):
            e2e["legs"]["moc100_geometric"] = geometric["e2e"]
    if not args.no_transport:
        transport = leg_transport(B, w, main["moc"])
        roofline["legs"]["transport"] = transport["roofline"]
        if "mules" in transport:
            roofline["legs"]["transport_mules"] = transport["mules"]["roofline"]
        if "implicit" in transport:
            roofline["legs"]["transport_implicit"] = transport["implicit"]["roofline"]

    # ---- CPU baselines on this box's host cores (rank 0, N = 1 only) -----------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        kind, run, how = cpu_arm()
        threads = host_threads()
        rate, secs = cpu_moc_rate(w, args.cpu_sample_cells, threads, run)
        rate1, secs1 = cpu_moc_rate(w, max(64, args.cpu_sample_cells // max(1, threads)), 1, run)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": kind,
               "sample": f"{args.cpu_sample_cells} cells of {w.name} (N={N}); {how}; {secs:.1f} s; per-cell rate, "
                         "no extrapolation needed (cells independent)",
               "single_thread_value": rate1, "single_thread_seconds": secs1, "legs": {}}
        if config1 is not None:
            # the FULL 1e5-cell configuration on the CPU
            r, s = cpu_moc_rate(w1, w1.ncells, threads, run)
            r1, s1 = cpu_moc_rate(w1, max(64, w1.ncells // (2 * threads)), 1, run)
            config1["cpu_baseline"] = {"value": r, "unit": UNIT, "cores": threads, "kind": kind,
                                       "sample": f"all {w1.ncells} cells of configs[0], {s:.1f} s", "seconds": s,
                                       "single_thread_value": r1}
            cpu["legs"]["config1_moc30"] = config1["cpu_baseline"]
        if qmom is not None:
            r, s, qkind, qhow = cpu_qmom_rate(wq, int(args.cpu_qmom_cells), threads)
            r1, s1, _, _ = cpu_qmom_rate(wq, max(1000, int(args.cpu_qmom_cells) // threads), 1)
            qmom["cpu_baseline"] = {"value": r, "unit": "cell-updates/s", "cores": threads, "kind": qkind,
                                    "sample": f"{int(args.cpu_qmom_cells)} cells of configs[1], reference-shaped: one "
                                              "wheeler_inversion per moment index (QMOM.C:231), pow/exp per kernel "
                                              f"evaluation; {qhow}; {s:.1f} s",
                                    "single_thread_value": r1}
            cpu["legs"]["qmom"] = qmom["cpu_baseline"]
        if replay is not None:
            r, s = cpu_moc_rate(wrm, 20_000, threads, run)
            replay["moc"]["cpu_baseline"] = {"value": r, "unit": UNIT, "cores": threads, "kind": kind,
                                             "sample": f"20000 cells of configs[4] (MOC 50 classes), {s:.1f} s",
                                             "ms_per_step_at_4e6_cells": 4e6 * wrm.N / r * 1e3}
            rq, sq, qk, _ = cpu_qmom_rate(wrq, 200_000, threads)
            replay["qmom"]["cpu_baseline"] = {"value": rq, "unit": "cell-updates/s", "cores": threads, "kind": qk,
                                              "sample": f"200000 cells of configs[4] (QMOM order 3), {sq:.1f} s",
                                              "ms_per_step_at_4e6_cells": 4e6 / rq * 1e3}
            cpu["legs"]["replay_moc50"] = replay["moc"]["cpu_baseline"]
            cpu["legs"]["replay_qmom3"] = replay["qmom"]["cpu_baseline"]
        if transport is not None:
            # the face loop of the transport update on one host core (the reference is serial per
            # rank), on a bounded sample: a 96^3 block, all fields
            from oracle import cpu_oracle as orc2

            sm_ = synthetic.block_mesh(96, 96, 96, velocity=(1.0, 0.5, 0.25))
            rs = np.random.default_rng(5)
            ps = rs.uniform(size=(N, 96 ** 3))
            Ss = rs.uniform(size=(N, 96 ** 3))
            t0 = time.perf_counter()
            orc2.transport_upwind_explicit(sm_["owner"], sm_["neighbour"], sm_["boundary_owner"], sm_["vcell"], ps,
                                           sm_["phi"], 1e-3, S=Ss)
            tsec = time.perf_counter() - t0
            transport["cpu_baseline"] = {"value": N * 96 ** 3 / tsec, "unit": "cell-field-updates/s", "cores": 1,
                                         "kind": "port", "sample": f"96^3 cells x {N} fields, one step, face loop of "
                                         f"oracle/pbe_oracle.c (OpenFOAM-shaped), {tsec:.2f} s"}
            cpu["legs"]["transport"] = transport["cpu_baseline"]

    if rank == 0:
        line = {
            "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": main["warmup"], "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{w.name}: MOC {N} classes x {cells} cells/GPU, CoulaloglouTavlaridesC "
                                   "coalescence + CoulaloglouTavlarides breakup + uniformBinaryBreakup (c01 "
                                   "constants), uniform class grid, eps log-uniform [1e-3,1], in-place zero-flux "
                                   "step + 6 fused global moments",
                       "cells_per_gpu": cells, "classes": N, "global_cells": world * cells,
                       "l2": "inputs (12.8 GB/GPU) exceed the 126 MB L2; no flush needed",
                       "parallelism": f"cells block-sharded over {world} GPU(s); NCCL all-reduce of 6 doubles/step"
                                      if world > 1 else "1 GPU",
                       "device": B.info["name"], "sm_count": B.info["sm_count"],
                       "other_legs": "config1 (configs[0]), qmom (configs[1]), strong (configs[3]), replay "
                                     "(configs[4]), transport, mom, moc100_geometric (configs[2] as worded: fixed-pivot extension): own keys of this line; their rooflines, CPU "
                                     "baselines and e2e numbers are repeated under roofline.legs, cpu_baseline.legs, "
                                     "e2e.legs"},
            "clocks": main["clocks"], "e2e": e2e, "gpu_launches": main["launches"], "roofline": roofline,
            "cpu_baseline": cpu, "strong": main["strong"], "config1": config1, "qmom": qmom, "replay": replay,
            "transport": transport, "mom": mom_leg, "moc100_geometric": geometric, "checks": main["checks"],
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        B.dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="moc100_ct")
    ap.add_argument("--cells", type=float, default=0, help="cells per GPU (default: the workload's)")
    ap.add_argument("--qmom-cells", type=float, default=0)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--roundtrip-cells", type=float, default=4_000_000,
                    help="cells of the e2e.roundtrip leg (all class fields over PCIe both ways)")
    ap.add_argument("--strong-cells", type=float, default=128_000_000, help="configs[3]: total cells of the strong-scaling line")
    ap.add_argument("--replay-cells", type=float, default=4_000_000)
    ap.add_argument("--replay-steps", type=int, default=10)
    ap.add_argument("--cpu-sample-cells", type=int, default=8000)
    ap.add_argument("--cpu-qmom-cells", type=float, default=1_000_000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-config1", action="store_true")
    ap.add_argument("--no-qmom", action="store_true")
    ap.add_argument("--no-replay", action="store_true")
    ap.add_argument("--no-transport", action="store_true")
    ap.add_argument("--no-mom", action="store_true")
    ap.add_argument("--no-geometric", action="store_true")
    ap.add_argument("--geo-cells", type=float, default=1_212_416, help="fixed-pivot (geometric grid) leg: cells per GPU")
    ap.add_argument("--geo-ratio", type=float, default=1.08, help="volume ratio of consecutive classes")
    ap.add_argument("--mom-cells", type=float, default=200000, help="methodOfMoments leg: cells per GPU")
    ap.add_argument("--halo-nccl", action="store_true", help="transport leg: NCCL send/recv halo exchange")
    ap.add_argument("--transport-n", type=int, default=160, help="transport leg: hex block of n^3 cells")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    _, world, _ = dist_env()
    if world > 1:
        args.gpus = world
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
