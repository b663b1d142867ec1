#!/usr/bin/env python
"""bench.py -- PBE source throughput of libpbe_b200.so on N B200s (one process per GPU).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU algorithm (oracle)

Metric (BASELINE.json): MOC cell-class-updates/s at 100 classes.  One *step* = one pass of the
hot path over this rank's cells: `pbe_moc_step` = breakup-rate evaluation + coalescence
birth/death double sums + breakup redistribution + limiter + zero-flux explicit update in place
(MOC::correct() with phi = 0, src/PBESystem/MOC.C:159-184, 259-272) with the fused global-moment
partial sums of BASELINE configs[3]; for N > 1 the 6 partial moments are all-reduced over NCCL on
a side stream (the only cross-rank traffic: cells are independent).

  * N = 1 workload = BASELINE configs[2]: 100 classes, 1.6e7 cells, Coulaloglou-Tavlarides
    coalescence + breakup with the c01 validation constants, synthetic eps / alpha fields.
  * N > 1: 1.6e7 cells PER GPU (weak scaling; N = 8 is configs[3]'s 1.28e8 cells).
  * `value`  : whole-job cell-class-updates/s, fields resident in HBM, CUDA-event time of exactly
               K steps between barriers, max over ranks.  Inputs (12.8 GB / rank) exceed L2.
  * `e2e`    : the same metric through the C-ABI host-buffer call an OpenFOAM adapter makes from
               MOC::correct() (`pbe_moc_source_host`: pinned host n -> S), host<->device copies
               inside the timed region.
  * `roofline`: the MOC kernel is FP64-pipe bound at 100 classes (66 flop/B); `roofline` reports
               the binding FP64 roof (peak = DFMA-chain micro-kernel measured in the same run --
               MEASURED_PEAKS.json has no FP64 figure) and `roofline.hbm` the HBM view against
               MEASURED_PEAKS.json.
  * `cpu_baseline`: oracle/ (C restatement of the reference loops, reference-shaped, all host
               threads) on a bounded slice of the same workload.  kind = "port": the reference
               itself needs OpenFOAM + Eigen3 + GSL and cannot be built here.
  * `qmom`   : the second half of the metric (QMOM 6 moments, BASELINE configs[1], 1e7 cells / GPU)
               with its own value / roofline, measured after the MOC region.
  * `mom`    : methodOfMoments (first slice): pbe_mom_step on 2e5 cells / GPU with
               the CT kernels at O(1) scales, FP64 roofline by the SURVEY flop convention.
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
#     of the tile loop and the epilogue): 18.9 per pair, 44 per class with 6 moments.  The FP64 pipe
#     retires one warp instruction per scheduler every 2 cycles whatever the opcode, so
#     instructions x 2 / DFMA-peak is the pipe utilisation (`roofline.fp64_pipe_frac`).
FP64_INSTR_PER_PAIR_CT = 14.3   # 13.25 per table slot x 2704 slots / 2500 valid pairs at N = 100 (8 % zero padding)
FP64_INSTR_PER_PAIR_TAB = 4.0
FP64_INSTR_PER_CLASS_CT = 42.3  # ncu r01w (arithmetic unchanged since): 39 975 FP64 warp instructions per cell in total at N = 100
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
    """dram bytes per launch from the committed ncu capture (profiles/r01_traffic.json), scaled by
    the cell count when the bench runs another size; None when there is no capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as fh:
            t = json.load(fh)[key]
        per_cell = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["cells"]
        return per_cell * cells
    except Exception:
        return None


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
# reference arm: the reference's CPU algorithm (oracle port), all host threads, bounded sample
# ==================================================================================================
def host_threads() -> int:
    """Host threads this process may use.  Not omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to every rank, which would time the reference arm on one core."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference_rate(w, sample_cells: int, nthreads: int, repeats: int = 1):
    """cell-class-updates/s of the reference-shaped CPU loops (oracle/pbe_oracle.c) on a slice."""
    from openpbe_b200 import synthetic
    from oracle import cpu_oracle as orc

    f = synthetic.moc_fields(w, ncells=sample_cells)
    sp = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    best = float("inf")
    for _ in range(repeats):
        t = time.perf_counter()
        orc.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt, nthreads=nthreads)
        best = min(best, time.perf_counter() - t)
    return sample_cells * w.N / best, best


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    from openpbe_b200 import synthetic
    from oracle import cpu_oracle as orc

    orc.build()
    w = synthetic.workload(args.workload)
    threads = host_threads()
    sample = args.cpu_sample_cells
    f = synthetic.moc_fields(w, ncells=sample)
    sp = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)

    def step():
        orc.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt, nthreads=threads)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = sample * w.N / dt
    desc = (f"{sample} cells of {w.name} (N={w.N}) per step, oracle/pbe_oracle.c reference-shaped loops "
            f"(MOC.C:92-157 order, pow/exp per ordered pair), OpenMP over cells")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{w.name}: MOC {w.N} classes, CT coalescence + CT breakup (c01 constants)",
                   "cells_per_step": sample, "note": "CPU rate is per-cell; the sample bounds run time"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ==================================================================================================
# this repo's arm
# ==================================================================================================
def run_b200(args):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from openpbe_b200 import capi, engine, sharding, synthetic

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libpbe_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL_DEBUG=VERSION/INFO prints its banner there
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "WARN"
        # ... and whatever the communicator set-up still writes to fd 1 goes to stderr instead
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            t = torch.zeros(1, device=dev)
            dist.all_reduce(t)  # creates the communicator now, inside the redirect
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)
    ctx = engine.Context(local)
    info = ctx.info()

    w = synthetic.workload(args.workload)
    cells = int(args.cells) if args.cells else w.ncells  # per GPU (weak scaling)
    N = w.N
    nmom = 6
    cell0 = rank * cells

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def allmax(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- inputs: generated chunk-wise on the host, resident copies in HBM; the first e2e_cells
    # cells also stay in pinned host memory as the e2e leg's buffers -------------------------------
    t_gen = time.perf_counter()
    e2e_cells = cells
    if args.no_e2e:
        e2e_cells = 0
    else:
        try:
            with open("/proc/meminfo") as fh:
                avail = [int(ln.split()[1]) * 1024 for ln in fh if ln.startswith("MemAvailable")][0]
        except Exception:
            avail = 64 << 30
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
        budget = int(0.5 * avail / max(1, local_world))
        if 2 * N * cells * 8 > budget:  # n + S pinned per rank
            e2e_cells = max(1, budget // (2 * N * 8))
    n_host = torch.empty((N, max(1, e2e_cells)), dtype=torch.float64, pin_memory=True)
    S_host = torch.empty((N, max(1, e2e_cells)), dtype=torch.float64, pin_memory=True)
    small_h = [torch.empty(max(1, e2e_cells), dtype=torch.float64, pin_memory=True) for _ in range(3)]
    eps_h, rho_h, nu_h = small_h
    n0 = torch.empty((N, cells), dtype=torch.float64, device=dev)
    eps, rho, nu, alpha, vcell = (torch.empty(cells, dtype=torch.float64, device=dev) for _ in range(5))
    chunk = min(cells, 2_000_000)
    st_n = np.empty((N, chunk))
    st_s = [np.empty(chunk) for _ in range(5)]
    for c0 in range(0, cells, chunk):
        nb = min(chunk, cells - c0)
        fill_moc_host(w, nb, cell0 + c0, st_n, *st_s)
        n0[:, c0:c0 + nb].copy_(torch.from_numpy(st_n[:, :nb]))
        for dst, src in zip((eps, rho, nu, alpha, vcell), st_s):
            dst[c0:c0 + nb].copy_(torch.from_numpy(src[:nb]))
        if c0 < e2e_cells:
            k = min(nb, e2e_cells - c0)
            n_host[:, c0:c0 + k].copy_(torch.from_numpy(st_n[:, :k]))
            for dst, src in zip(small_h, st_s[:3]):
                dst[c0:c0 + k].copy_(torch.from_numpy(src[:k]))
    del st_n, st_s
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t_gen

    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, N, w.xi1, spec, w.relax)
    n = n0.clone()
    moments = torch.zeros(nmom, dtype=torch.float64, device=dev)
    side = torch.cuda.Stream(device=dev)
    red_bufs = []

    def step():
        moments.zero_()
        moc.step(n, eps, rho, nu, w.dt, vcell=vcell, moments=moments)
        if world > 1:
            # global moments sum_c V_c sum_i xi_i^k n_i: 48 B, latency-bound, off the kernel stream
            buf = moments.clone()
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                dist.all_reduce(buf, op=dist.ReduceOp.SUM)
            red_bufs.append(buf)

    fp64_tf, _ = ctx.fp64_peak_tflops()

    warmup = max(args.warmup, 3)  # timing rule: at least 3 untimed steps
    for _ in range(warmup):
        step()
    red_bufs.clear()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    launches0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.current_stream(dev).wait_stream(side)
    barrier()
    tw1 = time.perf_counter()
    launches = ctx.launch_count() - launches0
    ms_local = e0.elapsed_time(e1)
    ms = allmax(ms_local)
    clocks = sampler.stop(tw0, tw1) if rank == 0 else None
    ms_per_step = ms / args.steps
    value = world * cells * N / (ms_per_step * 1e-3)
    gmom = red_bufs[-1].cpu().tolist() if red_bufs else moments.cpu().tolist()
    # realisability after the replayed steps: n >= 0 everywhere (MOC limiter, MOC.C:87-91)
    n_min = float(n.min().item())

    # ---- roofline of the dominant (only) kernel: one launch per step ---------------------------
    hbm_peak, hbm_src = load_peaks()
    kern_ms = ms_local / args.steps
    flops_cell = moc_flops_per_cell(w)
    bytes_cell = moc_bytes_per_cell(N) + 8.0  # + vcell for the moment sums
    tf = cells * flops_cell / (kern_ms * 1e-3) / 1e12
    gbs = cells * bytes_cell / (kern_ms * 1e-3) / 1e9
    fp64_nominal = info["sm_count"] * 64 * 2 * 1.965e9 / 1e12
    kinfo = moc.kernel()
    instr_cell = moc_fp64_instr_per_cell(w)
    pipe_frac = cells * instr_cell * 2.0 / (kern_ms * 1e-3) / 1e12 / fp64_tf
    roofline = {
        "bound": "fp64", "achieved": tf, "peak": fp64_tf, "unit": "TFLOP/s", "frac": tf / fp64_tf,
        "traffic": load_traffic("moc_tmem_kernel", cells) if (kinfo["kernel"] == "tmem" and N == 100) else None,
        "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum, profiles/r01_traffic.json "
                          "(captured at 1.6e7 cells; scaled per cell for other sizes)",
        "algorithmic_bytes": cells * bytes_cell, "kernel_ms": kern_ms,
        "kernel": ("moc_tmem_kernel<CT,CT>" if kinfo["kernel"] == "tmem" else "moc_source_kernel<CT,CT,128>")
                  + f" ({kinfo['cells_per_cta']} cells/CTA, {kinfo['ctas_per_sm']} CTA/SM)",
        "peak_source": "measured in this run: register-resident DFMA chains (pbe_measure_fp64_peak); "
                       f"nominal {fp64_nominal:.1f} TFLOP/s at 1965 MHz",
        "flops_per_cell": flops_cell,
        "flops_convention": "SURVEY.md 8(d): add/mul=1, FMA=2, exp=32: 40 + 41 P(N) + 40 N per cell",
        "fp64_pipe_frac": pipe_frac, "fp64_pipe_instr_per_cell": instr_cell,
        "fp64_pipe_note": "SASS-counted FP64-pipe instructions x 2 / DFMA peak = share of FP64 issue slots used",
        "hbm": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                "bytes_per_cell": bytes_cell, "peak_source": hbm_src},
    }

    # ---- e2e: host buffers through the C ABI (what the OpenFOAM adapter calls) -------------------
    e2e = None
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        moc.source_host(n_host, eps_h, rho_h, nu_h, w.dt, out=S_host)  # warm-up: staging buffers
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            moc.source_host(n_host, eps_h, rho_h, nu_h, w.dt, out=S_host)
        torch.cuda.synchronize()
        t_e2e = allmax((time.perf_counter() - t0) / e2e_steps)
        barrier()
        h2d = (N + 3) * e2e_cells * 8
        d2h = N * e2e_cells * 8
        e2e = {"value": world * e2e_cells * N / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": t_e2e * 1e3, "steps": e2e_steps,
               "call": "pbe_moc_source_host (pinned host n,eps,rho,nu -> S), synchronous",
               "pinned": True, "pcie_gbs": (h2d + d2h) / t_e2e / 1e9, "cells_per_gpu": e2e_cells,
               "note": "full workload" if e2e_cells == cells else
                       "host RAM bounds the pinned buffers: first e2e_cells cells of the workload per GPU"}
        # cheap end-to-end sanity: S from the host path equals S from the resident path on a slice
        chk = min(e2e_cells, 4096)
        S_dev = moc.source(n0[:, :chk].contiguous(), eps[:chk].contiguous(), rho[:chk].contiguous(),
                           nu[:chk].contiguous(), w.dt)
        if not torch.equal(S_dev.cpu(), S_host[:, :chk]):
            raise SystemExit("bench.py: host-buffer path and resident path disagree")

    # ---- QMOM half of the metric (configs[1]) -------------------------------------------------
    qmom = None
    if not args.no_qmom:
        wq = synthetic.workload("qmom3_ct")
        qc = int(args.qmom_cells) if args.qmom_cells else wq.ncells
        mh = torch.empty((2 * wq.N, qc), dtype=torch.float64)
        fq_small = [torch.empty(qc, dtype=torch.float64) for _ in range(3)]
        blk = 1_000_000
        for b0 in range(0, qc, blk):
            nb = min(blk, qc - b0)
            fq = synthetic.qmom_fields(wq, offset=rank * qc + b0, ncells=nb)
            mh.numpy()[:, b0:b0 + nb] = fq["m"]
            for t, k in zip(fq_small, ("eps", "rho", "nu")):
                t.numpy()[b0:b0 + nb] = fq[k]
        m0 = mh.to(dev)
        qe, qr, qn = (t.to(dev) for t in fq_small)
        qspec = engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs)
        q = engine.QMOM(ctx, wq.N, qspec)
        Sq = torch.empty_like(m0)
        dq = torch.empty(qc, dtype=torch.float64, device=dev)
        mq = m0.clone()
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

        def qstep():
            q.step(mq, qe, qr, qn, wq.dt, S=Sq, d=dq)

        # every step starts from the synthetic state (restored outside the timed region): with
        # quadrature weights of 1e5..1e7 the explicit update leaves most moments on the SMALL clamp
        # after one step (QMOM.C:117-123), and later steps would time the degenerate-cell path
        for _ in range(max(3, args.warmup)):
            mq.copy_(m0)
            qstep()
        barrier()
        tot = 0.0
        qsteps = max(3, args.steps)
        for _ in range(qsteps):
            mq.copy_(m0)
            flush.zero_()
            e0.record()
            qstep()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        qms = allmax(tot / qsteps)
        qbytes = 8.0 * (6 + 6 + 6 + 3 + 1)  # read 6 m + eps,rho,nu; write 6 m, 6 S, d
        qgbs = qc * qbytes / (qms * 1e-3) / 1e9
        qtf = qc * FLOPS_QMOM3_CT / (qms * 1e-3) / 1e12
        qmom = {"metric": "qmom6_cell_updates_per_s", "value": world * qc / (qms * 1e-3), "unit": "cell-updates/s",
                "ms_per_step": qms, "cells_per_gpu": qc,
                "workload": "qmom3_ct: Wheeler inversion + 6 moment sources + scrub/update/clamp/d, CT kernels",
                "l2": "flushed between timed launches (256 MB memset)",
                "roofline": {"bound": "hbm", "achieved": qgbs, "peak": hbm_peak, "unit": "GB/s",
                             "frac": qgbs / hbm_peak, "bytes_per_cell": qbytes, "traffic": load_traffic("qmom_kernel", qc),
                             "algorithmic_bytes": qc * qbytes,
                             "fp64_tflops": qtf, "fp64_frac": qtf / fp64_tf}}
        if not args.no_e2e:
            # the same metric through the host-buffer entry the adapter's QMOM::correct() calls
            # (pinned host m, eps, rho, nu -> S), copies inside the timed region
            mp = mh.pin_memory()
            sp = [t.pin_memory() for t in fq_small]
            Sh = torch.empty((2 * wq.N, qc), dtype=torch.float64, pin_memory=True)
            q.source_host(mp, *sp, out=Sh)
            barrier()
            t0 = time.perf_counter()
            for _ in range(3):
                q.source_host(mp, *sp, out=Sh)
            torch.cuda.synchronize()
            tq = allmax((time.perf_counter() - t0) / 3)
            qmom["e2e"] = {"value": world * qc / tq, "unit": "cell-updates/s", "ms_per_step": tq * 1e3,
                           "h2d_bytes_per_step": (2 * wq.N + 3) * qc * 8, "d2h_bytes_per_step": 2 * wq.N * qc * 8,
                           "call": "pbe_qmom_source_host (pinned host m,eps,rho,nu -> S), synchronous"}
            del mp, sp, Sh
        del m0, mq, Sq, flush

    # ---- methodOfMoments (first slice) -----------------------------------------
    mom_leg = None
    if not args.no_mom:
        mc = int(args.mom_cells)
        rng = np.random.default_rng(20240611 + rank)
        gm0, gk, gth = rng.uniform(0.5, 3.0, mc), rng.uniform(1.5, 4.0, mc), rng.uniform(0.15, 0.6, mc)
        mm0 = torch.from_numpy(np.stack([gm0, gm0 * gk * gth, gm0 * gk * (gk + 1) * gth ** 2])).to(dev)
        me, mr, mn = (torch.from_numpy(rng.uniform(lo, hi, mc)).to(dev)
                      for lo, hi in ((0.5, 2.0), (0.8, 1.2), (0.5, 1.5)))
        ct1 = (1.0, 1.0, 0.0, 1.0)  # O(1) scales: where integrate()'s epsabs 1e-7 means something
        mom = engine.MOM(ctx, engine.make_spec("CoulaloglouTavlaridesC", ct1, "CoulaloglouTavlarides", ct1))
        mwork, md = mm0.clone(), torch.empty(mc, dtype=torch.float64, device=dev)
        for _ in range(3):
            mwork.copy_(mm0)
            mom.step(mwork, me, mr, mn, 1e-3, d=md)
        barrier()
        tot, msteps = 0.0, min(5, max(3, args.steps))
        for _ in range(msteps):
            mwork.copy_(mm0)
            e0.record()
            mom.step(mwork, me, mr, mn, 1e-3, d=md)
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        mms = allmax(tot / msteps)
        nq = 15 * (12 + 20)
        # convention of SURVEY 8(d): exp = 32, sqrt / div = 1, FMA = 2; per unordered node pair of the CT
        # kernel 4 add + 7 mul + sqrt + div + exp + 2 FMA = 49; per node the gamma pdf (log, exp, 4) = 68
        mflops = 49.0 * (nq * (nq + 1) / 2) + 68.0 * nq
        mtf = mc * mflops / (mms * 1e-3) / 1e12
        mom_leg = {"metric": "mom3_cell_updates_per_s", "value": world * mc / (mms * 1e-3), "unit": "cell-updates/s",
                   "ms_per_step": mms, "cells_per_gpu": mc, "quadrature_nodes": nq,
                   "workload": "methodOfMoments: gamma closure + coalescence / breakup source integrals (CT kernels, "
                               "O(1) scales) on the fixed graded Gauss-Kronrod rule + scrub/update/clamp/d",
                   "l2": "compute-bound: 56 B per cell of HBM traffic, no flush needed",
                   "roofline": {"bound": "fp64", "achieved": mtf, "peak": fp64_tf, "unit": "TFLOP/s",
                                "frac": mtf / fp64_tf, "flops_per_cell": mflops, "traffic": None}}
        del mm0, mwork

    # ---- transport of the class fields (first slice) --------------------------
    transport = None
    if not args.no_transport:
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
        l0 = ctx.launch_count()

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
        barrier()
        tsteps = max(3, args.steps)
        e0.record()
        for _ in range(tsteps):
            tstep()
        e1.record()
        torch.cuda.synchronize()
        tms = allmax(e0.elapsed_time(e1) / tsteps)
        # psi and S in, psi out per cell-field; V, 6 x (neighbour, face) labels and 3 face fluxes per cell
        tbytes = tc * (24.0 * tK + 8 + 6 * 8 + 3 * 8)
        tgbs = tbytes / (tms * 1e-3) / 1e9
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
                     "roofline": {"bound": "hbm", "achieved": tgbs, "peak": hbm_peak, "unit": "GB/s",
                                  "frac": tgbs / hbm_peak, "algorithmic_bytes": tbytes,
                                  "traffic": load_traffic("transport_upwind_kernel", tc)}}
        if not args.no_e2e and world == 1:
            # MOC::correct() with solveWithMULES through the host-buffer entry the adapter calls
            # (pinned host class fields in, updated fields + d out; sources + transport on the device)
            ft = synthetic.moc_fields(w, ncells=min(tc, 200_000))
            rep = (tc + ft["n"].shape[1] - 1) // ft["n"].shape[1]
            nh = torch.from_numpy(np.tile(ft["n"], (1, rep))[:, :tc].copy()).pin_memory()
            sm = [torch.from_numpy(np.tile(ft[k], rep)[:tc].copy()).pin_memory() for k in ("eps", "rho", "nu")]
            ph = torch.from_numpy(tm["phi"]).pin_memory()
            tmesh.moc_correct_host(moc, nh, *sm, ph, w.dt)
            t0 = time.perf_counter()
            for _ in range(2):
                dh = tmesh.moc_correct_host(moc, nh, *sm, ph, w.dt)
            te = (time.perf_counter() - t0) / 2
            transport["e2e"] = {"value": tc * tK / te, "unit": "cell-class-updates/s", "ms_per_step": te * 1e3,
                                "h2d_bytes_per_step": (tK + 3) * tc * 8 + tm["phi"].size * 8,
                                "d2h_bytes_per_step": (tK + 1) * tc * 8,
                                "call": "pbe_moc_correct_host (pinned host n, eps, rho, nu, phi -> n, d): "
                                        "sources + upwind transport update, synchronous",
                                "d_mean": float(dh.mean())}
            del nh, sm, ph
        del tpsi, tS, tout, tphi, tmesh

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) -----------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import cpu_oracle as orc

        orc.build()
        threads = host_threads()
        rate, secs = cpu_reference_rate(w, args.cpu_sample_cells, threads)
        rate1, secs1 = cpu_reference_rate(w, max(64, args.cpu_sample_cells // max(1, threads)), 1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{args.cpu_sample_cells} cells of {w.name} (N={N}), reference-shaped loops "
                         f"(oracle/pbe_oracle.c, OpenMP over cells), {secs:.1f} s; per-cell rate, "
                         "no extrapolation needed (cells independent)",
               "single_thread_value": rate1, "single_thread_seconds": secs1}

    if cpu is not None and transport is not None:
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

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{w.name}: MOC {N} classes x {cells} cells/GPU, CoulaloglouTavlaridesC "
                                   "coalescence + CoulaloglouTavlarides breakup + uniformBinaryBreakup (c01 "
                                   "constants), uniform class grid, eps log-uniform [1e-3,1], in-place zero-flux "
                                   "step + 6 fused global moments",
                       "cells_per_gpu": cells, "classes": N, "global_cells": world * cells,
                       "l2": "inputs (12.8 GB/GPU) exceed the 126 MB L2; no flush needed",
                       "parallelism": f"cells block-sharded over {world} GPU(s); NCCL all-reduce of 6 doubles/step"
                                      if world > 1 else "1 GPU",
                       "device": info["name"], "sm_count": info["sm_count"]},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "cpu_baseline": cpu, "qmom": qmom, "transport": transport, "mom": mom_leg,
            "checks": {"n_min_after_steps": n_min, "global_moments": gmom, "input_fill_s": t_gen},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
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
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-sample-cells", type=int, default=8000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-qmom", action="store_true")
    ap.add_argument("--no-transport", action="store_true")
    ap.add_argument("--no-mom", action="store_true")
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
