#!/usr/bin/env python
"""Benchmark of the hot path: fused RK-stage kernels for 3-D Burgers (WENO5) and 3-D heat.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one time step = 3 fused RK stages over the whole field (one pass of the hot path).
Default workload `burgers3d_weak512`: BASELINE.json config 5's weak-scaling case, 512x512x(512*N)
double, WENO5-JS + SSP-RK3 with MultiGPU/Burgers3d_Baseline semantics, one z-slab of 512 planes
per GPU (at N=1 simply 512^3).  Other workloads: burgers3d_256 (C3), heat3d_256 (C2),
heat3d_strong1024 / burgers3d_strong1024 (C4/C5 strong scaling), heat2d_512 (C1).

Prints ONE JSON line (rank 0).  value = cell-updates of all ranks / device time (max over ranks);
e2e = same metric through the solver API with HOST buffers (upload + S steps + download);
roofline = algorithmic HBM bytes of the stage kernel / its measured launch time vs the measured
copy bandwidth; cpu_baseline = the reference's CPU path (C transcription of the Matlab prototypes)
on a bounded sample of the same workload.  With the default workload the line also carries
`also`: the heat workload of the same shape and, on more than one GPU, the two 1024^3 strong-
scaling workloads, each a full record with its own roofline, measured in the same process after
the default workload's timed region; and `halo`: the 3-plane exchange of one stage timed alone.

`--impl reference` runs the reference's own CPU implementation of the path on the host cores:
W + K real time steps of oracle/matlab_port.c (C transcription of heat3d.m / LFWENO5FDM3d.m,
north_star's "reference CPU path"; no Matlab/Octave on this image) on a bounded sample of the
workload, all cores, with the OpenMP restatement of the CUDA programs timed beside it.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (target, nx, ny, nz_per_gpu or total, scaling, lengths, coef, e2e steps)
    "burgers3d_weak512": dict(target="burgers3d_mgpu", n=(512, 512, 512), scaling="weak", L=(2.0, 2.0, 2.0), coef=0.3, e2e_steps=20),
    "burgers3d_256": dict(target="burgers3d_mgpu", n=(256, 256, 256), scaling="weak", L=(2.0, 2.0, 2.0), coef=0.3, e2e_steps=43),
    "burgers3d_sgpu_256": dict(target="burgers3d_sgpu", n=(256, 256, 256), scaling="replicas", L=(2.0, 2.0, 2.0), coef=0.3, e2e_steps=43),
    "burgers3d_strong1024": dict(target="burgers3d_mgpu", n=(1024, 1024, 1024), scaling="strong", L=(2.0, 2.0, 2.0), coef=0.3, e2e_steps=20),
    "heat3d_256": dict(target="heat3d_mgpu", n=(256, 256, 256), scaling="weak", L=(2.0, 2.0, 2.0), coef=1.0, e2e_steps=100),
    "heat3d_weak512": dict(target="heat3d_mgpu", n=(512, 512, 512), scaling="weak", L=(2.0, 2.0, 2.0), coef=1.0, e2e_steps=50),
    "heat3d_strong1024": dict(target="heat3d_mgpu", n=(1024, 1024, 1024), scaling="strong", L=(2.0, 2.0, 2.0), coef=1.0, e2e_steps=50),
    "heat2d_512": dict(target="heat2d_sgpu", n=(512, 512, 1), scaling="replicas", L=(10, 10, 1), coef=1.0, e2e_steps=1000),
    # the two y-slab programs (MultiGPU/Burgers2d_Baseline, MultiGPU/Diffusion2d_Baseline); n = (nx, rows per GPU, 1)
    "burgers2d_weak8192": dict(target="burgers2d_mgpu", n=(8192, 8192, 1), scaling="weak", L=(2.0, 2.0, 1), coef=0.4, e2e_steps=50),
    "heat2d_weak8192": dict(target="heat2d_mgpu", n=(8192, 8192, 1), scaling="weak", L=(2.0, 2.0, 1), coef=1.0, e2e_steps=100),
    "heat3d_sgpu_256": dict(target="heat3d_sgpu_baseline", n=(256, 256, 256), scaling="replicas", L=(10, 10, 10), coef=1.0, e2e_steps=100),
}
BYTES_PER_CELL_STEP = {8: 64.0, 4: 32.0}     # 16 + 24 + 24, per time step
DEFAULT_WORKLOAD = "burgers3d_weak512"


def is_two_d(w):
    return w["target"] in ("heat2d_mgpu", "burgers2d_mgpu")


def metric_name(w, precision):
    """one string for both arms (the driver pairs the repo arm with the reference arm by it)"""
    dim = "2D" if is_two_d(w) or w["target"] == "heat2d_sgpu" else "3D"
    return "Mcell-updates/s (%s %s, SSP-RK3, %s)" % (dim, "WENO5 Burgers" if "burgers" in w["target"] else "heat", precision)


def geometry(w, N):
    nx, ny, nz0 = w["n"]
    two_d = is_two_d(w)
    if two_d:
        ny = ny * N if w["scaling"] == "weak" else ny
    nz_total = nz0 * N if w["scaling"] == "weak" and not two_d else nz0
    if w["scaling"] == "replicas":
        nz_total = nz0
    lz = w["L"][2] * (N if w["scaling"] == "weak" else 1)
    ly = w["L"][1] * (N if two_d and w["scaling"] == "weak" else 1)
    return nx, ny, nz_total, ly, lz


def make_config(name, w, N, precision, halo):
    esz = 8 if precision == "f64" else 4
    nx, ny, nz_total, _, _ = geometry(w, N)
    two_d = is_two_d(w)
    nranks = N if w["scaling"] != "replicas" else 1
    cfg = {"workload": name, "target": w["target"], "grid": [nx, ny, nz_total], "precision": precision,
           "scheme": "WENO5-JS + Lax-Friedrichs + SSP-RK3" if "burgers" in w["target"] else "O4 %s Laplacian + SSP-RK3" % ("9-pt" if two_d else "13-pt"),
           "decomposition": ("%s-slab x%d" % ("y" if two_d else "z", N)) if N > 1 else "single GPU",
           "l2_policy": "no flush: working set 3 x %.2f GiB per GPU >> 126 MB L2" %
                        ((nx * (ny / max(N, 1) + 6) if two_d else nx * ny * (nz_total / max(N, 1) + 6)) * esz / 2**30),
           "halo": halo if nranks > 1 else "none"}
    if w["target"] == "heat2d_sgpu":
        cfg["l2_policy"] = "L2-resident by nature: 3 x 2 MiB field (launch-latency bound, SURVEY 8a2)"
    return cfg


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_profile(workload):
    """Per-launch numbers of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json,
    written by tools/ncu_traffic.py).  A capture describes ONE build of the kernel: it records the hash of the SASS of
    the instantiations it profiled (a regular expression over their mangled names), and is used only while the built library still carries the same hash
    (multigpu_advectiondiffusion_b200/kernel_hashes.json, written by the Makefile) -- otherwise the instruction
    counts and DRAM bytes would silently describe a kernel that no longer exists."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None, "no capture"
    try:
        d = json.load(open(p))
    except Exception:
        return None, None, "unreadable capture"
    prof = d.get(workload)
    if prof is None:
        return None, d.get("fp64_peak"), "no capture for this workload"
    want, fam = prof.get("sass_md5"), prof.get("kernel_family")
    have = None
    try:
        import hashlib
        built = json.load(open(os.path.join(ROOT, "multigpu_advectiondiffusion_b200", "kernel_hashes.json")))
        import re
        items = sorted((k, v) for k, v in built.items() if fam and re.search(fam, k))  # = tools/sass_hashes.py family_hash
        have = hashlib.md5(json.dumps(items).encode()).hexdigest() if items else None
    except Exception:
        pass
    if not want or want != have:
        return None, d.get("fp64_peak"), "capture is stale: it was taken on a different build of the kernel (%s)" % prof.get("source", "?")
    return prof, d.get("fp64_peak"), "current"


class ClockSampler:
    """SM clock, power and throttle reasons every 50 ms while the timed region runs (the quantities of the
    nvidia-smi clocks line in B200_PROFILING.md, read through NVML in a thread of this process).  NVML is
    initialised in __init__, i.e. BEFORE the warm-up: starting an `nvidia-smi -lms` process right before
    the timed region put its multi-second start-up on a multi-GPU box inside the region and stalled the
    GPUs it enumerates (4 GPUs, heat 512^3: 1.53 ms/step without a sampler, 16.6 ms/step with one)."""

    def __init__(self, index):
        self.index = index
        self.rows, self.h, self.nv, self.thread = [], None, None, None
        self.stop_flag = False
        if index is None:
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            # torch/CUDA ordinals follow CUDA_VISIBLE_DEVICES; NVML's do not
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                tok = vis.split(",")[index].strip()
                phys = int(tok) if tok.isdigit() else None
                if phys is None:
                    self.h = pynvml.nvmlDeviceGetHandleByUUID(tok)
            if self.h is None:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.sample()               # first query pays NVML's lazy set-up
            self.rows = []
        except Exception:
            self.h = None

    def sample(self):
        nv = self.nv
        try:
            reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM), nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM),
                          nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0, int(reasons)))

    def _loop(self):
        while not self.stop_flag:
            try:
                self.sample()
            except Exception:
                break
            time.sleep(0.05)

    def start(self):
        if self.h is None:
            return
        import threading
        self.rows = []
        self.stop_flag = False
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def summary(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            self.thread = None
            try:
                self.sample()           # one more at the end of the region
            except Exception:
                pass
        rows = self.rows
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        reasons = sorted(n for n, b in bits.items() if any(r[3] & b for r in rows))
        return {"sm_mhz": float(np.median([r[0] for r in rows])) if rows else None,
                "sm_max_mhz": float(max(r[1] for r in rows)) if rows else None,
                "power_w_max": float(max(r[2] for r in rows)) if rows else None,
                "reasons": reasons, "samples": len(rows), "source": "NVML (nvidia_ml_py), rank 0's GPU, 50 ms period"}


# ------------------------------------------------------------------------------------------------------
# The reference's CPU implementation of the path, timed on the host cores (cpu_baseline / --impl reference)
# ------------------------------------------------------------------------------------------------------
def cpu_reference(w, warmup, steps, step_seconds):
    """W + K real time steps of the reference's CPU path on a bounded sample of workload `w`.

    Primary arm ("reference CPU path" of north_star): oracle/matlab_port.c, the C transcription of
    Matlab_Prototipes/{DiffusionNd/heat3d.m + Laplace3d.m, DiffusionNd/heat2d.m + Laplace2d.m,
    InviscidBurgersNd/LFWENO5FDM3d.m + WENO5resAdv_{X,Y,Z}.m} with the fixed dt the CUDA drivers use.  Beside it:
    oracle/advdiff_oracle.c, the OpenMP restatement of the CUDA programs (the parity oracle), a few steps on the
    same sample.  2-D Burgers has no transcribed prototype: the restatement is the value there.
    The sample keeps the full x-y extent of the workload and as many z planes (rows in 2-D) as one time step of
    about `step_seconds` allows; throughput per cell does not depend on the plane count.  All host cores are
    used: the OpenMP thread count is set explicitly (torch.distributed.run exports OMP_NUM_THREADS=1)."""
    from oracle import oracle as orc
    cores = orc.use_all_cores()
    nx, ny, nz = w["n"]
    tgt = w["target"]
    L, M = orc.lib(), orc.mlb()
    dbl, dptr = C.c_double, C.POINTER(C.c_double)
    three_d = not (is_two_d(w) or tgt == "heat2d_sgpu")
    burgers = "burgers" in tgt

    M.mlb_burgers3d_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [dbl] * 5 + [C.c_int, C.c_int] + [dptr] * 3
    M.mlb_heat3d_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [dbl] * 4 + [C.c_int]
    M.mlb_heat2d_run.argtypes = [C.c_void_p, C.c_int, C.c_int] + [dbl] * 3 + [C.c_int]
    L.orc_burgers3d_mgpu_step_f64.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [dbl] * 4
    L.orc_heat3d_step_f64.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [dbl] * 4 + [C.c_int]

    def field(units):
        """a deterministic field of the sample's shape (values are immaterial: no data-dependent branches)"""
        if three_d:
            return orc.ic_burgers3d_mgpu(nx, ny, units - 6) if burgers else orc.ic_heat3d_mgpu(nx, ny, units - 6)
        if tgt == "heat2d_sgpu":
            return orc.ic_heat2d_sgpu(nx, ny)
        return orc.ic_burgers2d_mgpu(nx, units - 6) if burgers else orc.ic_heat2d_mgpu(nx, units - 6)

    dx = 2.0 / (nx - 1)

    def matlab_steps(u, n):
        """n time steps of the Matlab-prototype transcription, in place; returns seconds"""
        dt3, it3, t3 = dbl(), dbl(), dbl()
        t0 = time.perf_counter()
        if three_d and burgers:
            rc = M.mlb_burgers3d_run(u.ctypes.data, nx, ny, u.shape[0], dx, dx, dx, 0.3, 1e9, 0, n, C.byref(dt3), C.byref(it3), C.byref(t3))
        elif three_d:
            rc = M.mlb_heat3d_run(u.ctypes.data, nx, ny, u.shape[0], 1.0 / dx**2, 1.0 / dx**2, 1.0 / dx**2, 0.1 * dx * dx, n)
        else:
            rc = M.mlb_heat2d_run(u.ctypes.data, nx, u.shape[0], 1.0 / dx**2, 1.0 / dx**2, 0.2 * dx * dx, n)
        assert rc == 0
        return time.perf_counter() - t0

    def port_steps(u, n):
        """n time steps of the OpenMP restatement of the CUDA programs; returns seconds"""
        if three_d:
            uo, Lu = np.empty_like(u), np.zeros_like(u)
            t0 = time.perf_counter()
            for _ in range(n):
                if burgers:
                    L.orc_burgers3d_mgpu_step_f64(u.ctypes.data, uo.ctypes.data, Lu.ctypes.data, nx, ny, u.shape[0], dx, dx, dx, 0.3 * dx)
                else:
                    L.orc_heat3d_step_f64(u.ctypes.data, uo.ctypes.data, Lu.ctypes.data, nx, ny, u.shape[0], 1.0, 1.0, 1.0, 1e-6, 2)
            return time.perf_counter() - t0
        t0 = time.perf_counter()
        if tgt == "heat2d_sgpu":
            orc.heat2d(u, n)
        elif burgers:
            orc.burgers2d_mgpu(u, 1e9, W=2.0 * (u.shape[0] - 6) / nx, max_steps=n)
        else:
            orc.heat2d_mgpu(u, n, W=2.0 * (u.shape[0] - 6) / nx)
        return time.perf_counter() - t0

    has_matlab = three_d or not burgers
    primary = matlab_steps if has_matlab else port_steps
    # calibrate the sample: one step on a thin sample gives the rate; then size for `step_seconds` per step
    if tgt == "heat2d_sgpu":
        units, unit_cells = ny, nx
    else:
        unit_cells = nx * ny if three_d else nx
        full = (nz if three_d else ny) + 6
        thin = 20 if three_d else 256
        u = field(thin)
        primary(u, 1)
        t1 = primary(u, 1)
        units = int(min(max(thin, step_seconds * thin / t1), full))
    u = field(units)
    cells = unit_cells * u.shape[0]
    primary(u, max(1, warmup))
    t = primary(u, steps)
    value = cells * 3 * steps / t / 1e6
    shape = ("%dx%dx%d planes" % (nx, ny, u.shape[0])) if three_d else ("%d x %d rows" % (nx, u.shape[0]))
    src = {True: "oracle/matlab_port.c, the C transcription of the reference's Matlab prototype (%s), OpenMP" %
                 ("LFWENO5FDM3d.m + WENO5resAdv_X/Y/Z.m" if burgers else ("heat3d.m + Laplace3d.m" if three_d else "heat2d.m + Laplace2d.m")),
           False: "oracle/advdiff_oracle.c, the OpenMP restatement of MultiGPU/Burgers2d_Baseline (no 2-D Burgers prototype is transcribed)"}[has_matlab]
    out = {"value": value, "unit": "Mcell-updates/s", "cores": cores, "kind": "port",
           "sample": "%s of the workload, %d warm-up + %d timed time steps (3 RK stages each), %s" % (shape, max(1, warmup), steps, src),
           "seconds": t, "ms_per_step": 1e3 * t / steps, "steps": steps}
    if has_matlab:
        try:                # the second opinion must never cost the bench line
            v = field(units)
            port_steps(v, 1)
            n2 = max(1, min(3, steps))
            t2 = port_steps(v, n2)
            out["cuda_program_restatement"] = {"value": cells * 3 * n2 / t2 / 1e6, "unit": "Mcell-updates/s", "cores": cores,
                                               "sample": "same sample, %d time steps of oracle/advdiff_oracle.c (OpenMP restatement of the reference's CUDA program, the parity oracle)" % n2}
        except Exception as e:
            out["cuda_program_restatement"] = {"error": str(e)[:200]}
    return out


# ------------------------------------------------------------------------------------------------------
# one workload on the GPUs
# ------------------------------------------------------------------------------------------------------
def run_gpu_workload(name, ctx, steps, warmup, precision, halo_mode, with_e2e, with_cpu, sampler):
    import torch
    import multigpu_advectiondiffusion_b200 as adv
    dist, rank, world, dev = ctx["dist"], ctx["rank"], ctx["world"], ctx["dev"]
    w = WORKLOADS[name]
    N = world
    esz = 8 if precision == "f64" else 4
    nx, ny, nz_total, ly, lz = geometry(w, N)
    two_d = is_two_d(w)
    nranks = world if w["scaling"] != "replicas" else 1

    comm_id = None
    if nranks > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            adv._lib.check(adv.lib().adv_comm_unique_id(C.byref(buf)))
            idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm_id = bytes(idt.cpu().tolist())

    s = adv.Solver(w["target"], nx, ny, nz_total, lx=w["L"][0], ly=ly, lz=lz, coef=w["coef"],
                   precision=adv.F64 if esz == 8 else adv.F32, device=dev, nranks=nranks, rank=rank if nranks > 1 else 0,
                   comm_id=comm_id)
    if nranks > 1 and halo_mode == "peer":
        hs = [None] * nranks
        dist.all_gather_object(hs, s.peer_handle())
        s.peer_connect(hs[rank - 1] if rank > 0 else None, hs[rank + 1] if rank < nranks - 1 else None)
    config = make_config(name, w, N, precision, s.halo_mode)
    s.init_reference_ic_device()        # the reference's closed form, evaluated on the device (throughput run)
    cells_local = nx * (ny // nranks) if two_d else nx * ny * (nz_total // nranks)
    cells_job = cells_local * (world if w["scaling"] == "replicas" or nranks > 1 else 1)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    s.step(warmup)
    s.sync()
    barrier()
    if sampler is not None:
        sampler.start()
    l0 = s.kernel_launches
    t_wall0 = time.perf_counter()
    s.step(steps)
    s.sync()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = s.kernel_launches - l0
    ms = max_over_ranks(s.elapsed_ms())
    clocks = sampler.summary() if sampler is not None else None
    value = cells_job * 3 * steps / (ms * 1e-3) / 1e6

    # roofline of the dominant kernel (the fused stage kernel; one launch per stage at N=1)
    peak, peak_src = measured_peak()
    per_launch_bytes = BYTES_PER_CELL_STEP[esz] / 3.0 * cells_local
    launch_s = ms * 1e-3 / (3 * steps)
    achieved = per_launch_bytes / launch_s / 1e9
    prof, fp64_peak, prof_state = ncu_profile(name)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": prof["dram_bytes_per_launch_mean"] if prof else None, "traffic_capture": prof_state, "peak_source": peak_src,
                "kernel": ("burgers2d_stage_tile" if "burgers" in w["target"] else "heat2d_stage_rows") if two_d else
                          ("burgers3d_stage" if "burgers" in w["target"] else "heat3d_stage"),
                "algorithmic_bytes_per_launch": per_launch_bytes,
                "note": "FP64 WENO5 is bound by the FP64 pipe, not HBM (SURVEY F11); see DESIGN.md" if "burgers" in w["target"] and esz == 8 else ""}

    # FP64 WENO5 is bound by the FP64 pipe: FP64 thread-instructions per launch (ncu capture of THIS build of the
    # kernel, see ncu_profile) over the live launch time, against the measured DFMA peak (tools/fp64_peak.cu)
    fp64 = None
    if prof and fp64_peak and "burgers" in w["target"] and esz == 8:
        inst = float(np.mean(prof["warp_inst_per_launch_stage123"])) * prof["fp64_share_of_inst_stage2"] * 32.0
        fp64 = {"bound": "fp64-pipe", "achieved": inst / launch_s / 1e12, "peak": fp64_peak["fma_per_s"] / 1e12,
                "unit": "T FP64 instr/s", "frac": inst / launch_s / fp64_peak["fma_per_s"],
                "fp64_inst_per_cell_update": inst / cells_local,
                "source": "ncu warp-instruction count x FP64 share (profiles/traffic.json, same kernel SASS as this build) / live launch time; peak = measured DFMA rate"}
    elif "burgers" in w["target"] and esz == 8:
        fp64 = {"unavailable": prof_state}

    # the exchange of one stage alone (3 planes per neighbour), timed on the device
    halo = None
    if nranks > 1:
        per_nbr = 3 * nx * (1 if two_d else ny) * esz
        barrier()
        hms, hbytes = s.halo_probe(50)
        barrier()
        hms_max = max_over_ranks(hms)
        middle = 2 if nranks > 2 else 1
        halo = {"bytes_per_neighbour": per_nbr, "bytes_sent_per_gpu_per_exchange": per_nbr * middle,
                "exchange_us": hms_max * 1e3,
                "GBps_per_direction": per_nbr / (hms_max * 1e-3) / 1e9,
                "GBps_sent_per_gpu": per_nbr * middle / (hms_max * 1e-3) / 1e9,
                "nvlink_peak_GBps_per_direction": 900.0, "frac_of_nvlink": per_nbr / (hms_max * 1e-3) / 1e9 / 900.0,
                "mode": s.halo_mode, "exchanges_per_s_in_the_run": 3 * steps / (ms * 1e-3),
                "share_of_step_if_not_overlapped": 3 * hms_max / (ms / steps),
                "definition": "adv_halo_probe: 50 exchanges of the 3 boundary planes per neighbour with nothing else running, CUDA events on "
                              "the issuing stream, max over ranks; one exchange = every rank's sends landed and flagged.  In the time "
                              "loop the same stores are part of the boundary-slab stage kernels and overlap the interior kernel."}

    # end to end through the solver API with host buffers: upload + S steps + download per e2e step
    e2e = None
    if with_e2e:
        S = w["e2e_steps"]
        host = torch.empty(s.local_shape, dtype=torch.float64 if esz == 8 else torch.float32).pin_memory()
        hnp = host.numpy()
        s.download(hnp)
        reps = max(1, min(3, steps))
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            s.upload(hnp)
            s.step(S)
            s.download(hnp)
        barrier()
        te = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": cells_job * 3 * S * reps / te / 1e6, "unit": "Mcell-updates/s",
               "h2d_bytes_per_step": int(hnp.nbytes), "d2h_bytes_per_step": int(hnp.nbytes),
               "definition": "one e2e step = adv_upload(pinned host slab) + %d time steps + adv_download, as one run of the reference driver" % S,
               "steps_per_call": S, "calls": reps}
        # where the end-to-end time goes: the two host copies timed alone, every rank copying at the same moment
        # (as in the loop above), beside the device time of the S steps.  With N ranks on one host the copies share
        # the host's memory and PCIe root complexes, so their per-GPU rate falls as N grows while the kernels' does not.
        barrier()
        t0 = time.perf_counter()
        s.upload(hnp)
        barrier()
        t_up = max_over_ranks(time.perf_counter() - t0)
        t0 = time.perf_counter()
        s.download(hnp)
        barrier()
        t_down = max_over_ranks(time.perf_counter() - t0)
        t_comp = S * (ms / steps) * 1e-3
        e2e["breakdown"] = {"h2d_s": t_up, "d2h_s": t_down, "device_steps_s": t_comp, "sum_s": t_up + t_down + t_comp,
                            "measured_s_per_call": te / reps,
                            "h2d_GBps_per_gpu": hnp.nbytes / t_up / 1e9, "d2h_GBps_per_gpu": hnp.nbytes / t_down / 1e9,
                            "host_copy_GBps_all_gpus": (world if nranks > 1 or w["scaling"] == "replicas" else 1) * 2 * hnp.nbytes / (t_up + t_down) / 1e9,
                            "copy_share_of_call": (t_up + t_down) / (t_up + t_down + t_comp)}
        del host, hnp

    s.close()
    cb = None
    if with_cpu and rank == 0:
        cb = cpu_reference(w, warmup=1, steps=3, step_seconds=3.0)

    return {"metric": metric_name(w, precision),
            "value": value, "unit": "Mcell-updates/s", "n_gpus": N, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
            "dtype": precision, "data": "synthetic", "config": config,
            "hbm_gbs": achieved * (world if nranks > 1 or w["scaling"] == "replicas" else 1),
            "roofline": roofline, "roofline_fp64": fp64, "halo": halo, "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks, "wall_s_timed_region": t_wall}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="f64", choices=["f64", "f32"])
    ap.add_argument("--halo", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU halo exchange: fused NVLink peer stores (default) or NCCL send/recv")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="default workload only: skip the heat / strong-scaling records")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    N = args.gpus

    # ---------------------------------------------------------------- reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        # every step is real: W warm-up + K timed time steps on a sample sized for ~0.4 s per step
        cb = cpu_reference(w, warmup=args.warmup, steps=args.steps, step_seconds=0.4)
        v = cb["value"]
        line = {"impl": "reference", "metric": metric_name(w, args.precision),
                "value": v, "unit": "Mcell-updates/s", "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
                "dtype": args.precision, "data": "synthetic",
                "config": make_config(args.workload, w, N, args.precision, args.halo), "cpu_baseline": dict(cb),
                "e2e": {"value": v, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "reference arm = the reference's own CPU implementation of the path on this box's host cores: its Matlab "
                        "prototype, transcribed to C (no Matlab/Octave on this image), all cores; ms_per_step is one time step of the "
                        "bounded sample named in cpu_baseline.sample (throughput per cell does not depend on the plane count)"}
        print(json.dumps(line))
        return 0

    # ---------------------------------------------------------------- our arm (GPU)
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback exists)")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == N or world == 1, "launch with torchrun --nproc-per-node == --gpus"
    ctx = {"dist": dist, "rank": rank, "world": world, "dev": local_rank}

    # one sampler for the job (rank 0's GPU), NVML initialised before the warm-up
    sampler = ClockSampler(local_rank if rank == 0 and not os.environ.get("ADV_NO_SAMPLER") else None)
    line = run_gpu_workload(args.workload, ctx, args.steps, args.warmup, args.precision, args.halo,
                            with_e2e=not args.no_e2e, with_cpu=(N == 1 and not args.no_cpu_baseline), sampler=sampler)

    # the other half of BASELINE.json's metric and its strong-scaling configurations, measured in the same process
    # (same contract per record: W warm-up steps, K timed steps, device time, max over ranks)
    if args.workload == DEFAULT_WORKLOAD and not args.no_also:
        also = {}
        extra = [("heat3d_weak512", args.steps)]
        if world > 1:
            extra += [("burgers3d_strong1024", max(3, min(args.steps, 5))), ("heat3d_strong1024", max(3, min(args.steps, 10)))]
        for name, k in extra:
            try:
                rec = run_gpu_workload(name, ctx, k, 3, args.precision, args.halo, with_e2e=False, with_cpu=False, sampler=sampler)
                rec.pop("cpu_baseline", None)
                rec.pop("e2e", None)
                also[name] = rec
            except Exception as e:          # an extra record must never cost the headline line
                also[name] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        line["also"] = also

    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
