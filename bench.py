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
copy bandwidth; cpu_baseline = the CPU oracle (OpenMP port of the reference algorithm) on a
bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
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


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_profile(workload):
    """Per-launch numbers of the dominant kernel from the committed ncu --set full capture
    (profiles/traffic.json, written by the summary step in profiles/README.md)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return d.get(workload), d.get("fp64_peak")
        except Exception:
            pass
    return None, None


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
        self.stop_flag = False
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def summary(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
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


def cpu_sample(w, precision_bytes, target_seconds=12.0):
    """Times the CPU oracle (OpenMP, all host threads) on a bounded sample of the workload:
    the full x-y extent and as many z planes as ~target_seconds of CPU work allow."""
    from oracle import oracle as orc
    import ctypes as C
    nx, ny, _ = w["n"]
    L = orc.lib()
    dt_ = np.float64
    threads = orc.num_threads()

    def run(planes, steps):
        if w["target"].startswith("burgers"):
            u = orc.ic_burgers3d_mgpu(nx, ny, planes - 6)
            uo, Lu = np.empty_like(u), np.empty_like(u)
            dx = 2.0 / (nx - 1)
            f = L.orc_burgers3d_mgpu_step_f64
            f.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [C.c_double] * 4
            t0 = time.perf_counter()
            for _ in range(steps):
                f(u.ctypes.data, uo.ctypes.data, Lu.ctypes.data, nx, ny, planes, dx, dx, dx, 0.3 * dx)
            return time.perf_counter() - t0
        u = orc.ic_heat3d_mgpu(nx, ny, planes - 6)
        uo, Lu = np.empty_like(u), np.zeros_like(u)
        f = L.orc_heat3d_step_f64
        f.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [C.c_double] * 4 + [C.c_int]
        t0 = time.perf_counter()
        for _ in range(steps):
            f(u.ctypes.data, uo.ctypes.data, Lu.ctypes.data, nx, ny, planes, 1.0, 1.0, 1.0, 1e-6, 2)
        return time.perf_counter() - t0

    if w["target"] in ("heat2d_mgpu", "burgers2d_mgpu"):
        rows = 2048                                    # full row length, a bounded number of rows
        if w["target"] == "burgers2d_mgpu":
            u0, steps = orc.ic_burgers2d_mgpu(nx, rows, W=2.0 * rows / nx), 4
            fn = lambda: orc.burgers2d_mgpu(u0, 1e9, W=2.0 * rows / nx, max_steps=steps)
        else:
            u0, steps = orc.ic_heat2d_mgpu(nx, rows), 20
            fn = lambda: orc.heat2d_mgpu(u0, steps, W=2.0 * rows / nx)
        fn()
        t0 = time.perf_counter()
        fn()
        t = time.perf_counter() - t0
        return {"value": nx * rows * 3 * steps / t / 1e6, "unit": "Mcell-updates/s", "cores": threads, "kind": "port",
                "sample": "%d x %d rows of the workload, %d time steps, oracle/advdiff_oracle.c with OpenMP" % (nx, rows, steps), "seconds": t}
    if w["target"] == "heat2d_sgpu":
        u0 = orc.ic_heat2d_sgpu(nx, ny)
        steps = 300
        t0 = time.perf_counter()
        orc.heat2d(u0, steps)
        t = time.perf_counter() - t0
        return {"value": nx * ny * 3 * steps / t / 1e6, "unit": "Mcell-updates/s", "cores": threads, "kind": "port",
                "sample": "%dx%d, %d time steps, oracle/advdiff_oracle.c with OpenMP" % (nx, ny, steps), "seconds": t}
    def run_matlab_port(planes):
        """the same sample through oracle/matlab_port.c (C transcription of the Matlab prototypes, north_star's
        'reference CPU path'); reported beside the faster restatement of the CUDA programs"""
        if w["target"].startswith("burgers"):
            u = orc.ic_burgers3d_mgpu(nx, ny, planes - 6)
            dx = 2.0 / (nx - 1)
            t0 = time.perf_counter()
            orc.mlb_burgers3d(u, 1e9, CFL=0.3, dx=dx, dy=dx, dz=dx, max_steps=1)
            return time.perf_counter() - t0
        u = orc.ic_heat3d_mgpu(nx, ny, planes - 6)
        t0 = time.perf_counter()
        orc.mlb_heat3d(u, 1, L=2.0, W=2.0, H=2.0)
        return time.perf_counter() - t0

    planes = 14
    t = run(planes, 1)                                  # calibration (also warms the pages)
    rate = nx * ny * planes * 3 / t
    planes = int(min(max(14, target_seconds * rate / (nx * ny * 3)), 262))
    steps = 1
    t = min(run(planes, steps) for _ in range(2))
    cells = nx * ny * planes
    try:
        pm = min(planes, 40)
        tm = run_matlab_port(pm)
        matlab = {"value": nx * ny * pm * 3 / tm / 1e6, "unit": "Mcell-updates/s",
                  "sample": "%dx%dx%d, 1 time step, oracle/matlab_port.c (C transcription of heat3d.m / LFWENO5FDM3d.m) with OpenMP" % (nx, ny, pm)}
    except Exception as e:      # the second opinion must never cost the bench line
        matlab = {"error": str(e)[:200]}
    return {"value": cells * 3 * steps / t / 1e6, "unit": "Mcell-updates/s", "cores": threads, "kind": "port", "matlab_transcription": matlab,
            "sample": "%dx%dx%d planes of the workload, %d time step (3 RK stages), oracle/advdiff_oracle.c with OpenMP, best of 2"
                      % (nx, ny, planes, steps), "seconds": t}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="burgers3d_weak512", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="f64", choices=["f64", "f32"])
    ap.add_argument("--halo", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU halo exchange: fused NVLink peer stores (default) or NCCL send/recv")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    N = args.gpus
    esz = 8 if args.precision == "f64" else 4
    nx, ny, nz0 = w["n"]
    two_d = w["target"] in ("heat2d_mgpu", "burgers2d_mgpu")        # y-slabs: the split axis is y
    if two_d:
        ny = ny * N if w["scaling"] == "weak" else ny
    nz_total = nz0 * N if w["scaling"] == "weak" and not two_d else nz0
    if w["scaling"] == "replicas":
        nz_total = nz0
    lz = w["L"][2] * (N if w["scaling"] == "weak" else 1)
    ly = w["L"][1] * (N if two_d and w["scaling"] == "weak" else 1)
    config = {"workload": args.workload, "target": w["target"], "grid": [nx, ny, nz_total], "precision": args.precision,
              "scheme": "WENO5-JS + Lax-Friedrichs + SSP-RK3" if "burgers" in w["target"] else "O4 %s Laplacian + SSP-RK3" % ("9-pt" if two_d else "13-pt"),
              "decomposition": ("%s-slab x%d" % ("y" if two_d else "z", N)) if N > 1 else "single GPU",
              "l2_policy": "no flush: working set 3 x %.2f GiB per GPU >> 126 MB L2" %
                           ((nx * (ny / max(N, 1) + 6) if two_d else nx * ny * (nz_total / max(N, 1) + 6)) * esz / 2**30)}

    # ---------------------------------------------------------------- reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        vals = []
        cb = None
        for i in range(max(1, min(args.steps, 3))):
            cb = cpu_sample(w, esz, target_seconds=8.0)
            vals.append(cb["value"])
        v = max(vals)
        cb["value"] = v
        line = {"impl": "reference", "metric": "Mcell-updates/s (%s %s, SSP-RK3)" % ("2D" if two_d or w["target"] == "heat2d_sgpu" else "3D", "WENO5 Burgers" if "burgers" in w["target"] else "heat"),
                "value": v, "unit": "Mcell-updates/s", "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * cb["seconds"], "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
                "dtype": args.precision, "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": v, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "reference arm = the reference algorithm on the host cores (no Matlab/Octave and no MPI on this "
                        "image): OpenMP C port oracle/advdiff_oracle.c, pinned bit-exact to the reference's own kernels"}
        print(json.dumps(line))
        return 0

    # ---------------------------------------------------------------- our arm (GPU)
    import torch
    import multigpu_advectiondiffusion_b200 as adv
    import ctypes as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback exists)")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == N or world == 1, "launch with torchrun --nproc-per-node == --gpus"
    dev = local_rank

    comm_id = None
    nranks = world if w["scaling"] != "replicas" else 1
    if nranks > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            adv._lib.check(adv.lib().adv_comm_unique_id(C.byref(buf)))
            idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm_id = bytes(idt.cpu().tolist())

    if w["target"] == "heat2d_sgpu":
        config["l2_policy"] = "L2-resident by nature: 3 x 2 MiB field (launch-latency bound, SURVEY 8a2)"
    s = adv.Solver(w["target"], nx, ny, nz_total, lx=w["L"][0], ly=ly, lz=lz, coef=w["coef"],
                   precision=adv.F64 if esz == 8 else adv.F32, device=dev, nranks=nranks, rank=rank if nranks > 1 else 0,
                   comm_id=comm_id)
    if nranks > 1 and args.halo == "peer":
        hs = [None] * nranks
        dist.all_gather_object(hs, s.peer_handle())
        s.peer_connect(hs[rank - 1] if rank > 0 else None, hs[rank + 1] if rank < nranks - 1 else None)
    config["halo"] = s.halo_mode
    s.init_reference_ic()
    cells_local = nx * (ny // nranks) if two_d else nx * ny * (nz_total // nranks)
    cells_job = cells_local * (world if w["scaling"] == "replicas" or nranks > 1 else 1)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # one sampler for the job (rank 0's GPU), NVML initialised before the warm-up
    sampler = ClockSampler(dev if rank == 0 and not os.environ.get("ADV_NO_SAMPLER") else None)
    s.step(args.warmup)
    s.sync()
    barrier()
    sampler.start()
    l0 = s.kernel_launches
    t_wall0 = time.perf_counter()
    s.step(args.steps)
    s.sync()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = s.kernel_launches - l0
    ms = s.elapsed_ms()
    clocks = sampler.summary()
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = cells_job * 3 * args.steps / (ms * 1e-3) / 1e6

    # roofline of the dominant kernel (the fused stage kernel; one launch per stage at N=1)
    peak, peak_src = measured_peak()
    per_launch_bytes = BYTES_PER_CELL_STEP[esz] / 3.0 * cells_local
    launch_s = ms * 1e-3 / (3 * args.steps)
    achieved = per_launch_bytes / launch_s / 1e9
    prof, fp64_peak = ncu_profile(args.workload)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": prof["dram_bytes_per_launch_mean"] if prof else None, "peak_source": peak_src,
                "kernel": ("burgers2d_stage_tile" if "burgers" in w["target"] else "heat2d_stage_rows") if two_d else
                          ("burgers3d_stage" if "burgers" in w["target"] else "heat3d_stage"),
                "algorithmic_bytes_per_launch": per_launch_bytes,
                "note": "FP64 WENO5 is bound by the FP64 pipe, not HBM (SURVEY F11); see DESIGN.md" if "burgers" in w["target"] and esz == 8 else ""}

    # FP64 WENO5 is bound by the FP64 pipe: FP64 thread-instructions per launch (ncu capture of this workload)
    # over the live launch time, against the measured DFMA peak (tools/fp64_peak.cu)
    fp64 = None
    if prof and fp64_peak and "burgers" in w["target"] and esz == 8:
        inst = float(np.mean(prof["warp_inst_per_launch_stage123"])) * prof["fp64_share_of_inst_stage2"] * 32.0
        fp64 = {"bound": "fp64-pipe", "achieved": inst / launch_s / 1e12, "peak": fp64_peak["fma_per_s"] / 1e12,
                "unit": "T FP64 instr/s", "frac": inst / launch_s / fp64_peak["fma_per_s"],
                "fp64_inst_per_cell_update": inst / cells_local,
                "source": "ncu warp-instruction count x FP64 share (profiles/traffic.json) / live launch time; peak = measured DFMA rate"}

    # end to end through the solver API with host buffers: upload + S steps + download per e2e step
    e2e = None
    if not args.no_e2e:
        S = w["e2e_steps"]
        host = torch.empty(s.local_shape, dtype=torch.float64 if esz == 8 else torch.float32).pin_memory()
        hnp = host.numpy()
        s.download(hnp)
        reps = max(1, min(3, args.steps))
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            s.upload(hnp)
            s.step(S)
            s.download(hnp)
        barrier()
        te = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([te], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            te = float(t.item())
        e2e = {"value": cells_job * 3 * S * reps / te / 1e6, "unit": "Mcell-updates/s",
               "h2d_bytes_per_step": int(hnp.nbytes), "d2h_bytes_per_step": int(hnp.nbytes),
               "definition": "one e2e step = adv_upload(pinned host slab) + %d time steps + adv_download, as one run of the reference driver" % S,
               "steps_per_call": S, "calls": reps}

    halo = None
    if nranks > 1:
        per_nbr = 3 * nx * (1 if two_d else ny) * esz
        nb = 2 if 0 < rank < nranks - 1 else 1
        halo = {"bytes_per_gpu_per_stage": per_nbr * (2 if nranks > 2 else 1), "bytes_per_neighbour": per_nbr,
                "stages_per_s": 3 * args.steps / (ms * 1e-3),
                "sustained_GBps_per_gpu": per_nbr * (2 if nranks > 2 else 1) * 3 * args.steps / (ms * 1e-3) / 1e9,
                "nvlink_peak_GBps": 900.0, "mode": s.halo_mode,
                "note": "sustained rate over the whole step (exchange is overlapped with the interior kernel, not timed alone)"}
    cb = None
    if rank == 0 and N == 1 and not args.no_cpu_baseline:
        cb = cpu_sample(w, esz)

    s.close()
    if rank == 0:
        line = {"metric": "Mcell-updates/s (%s %s, SSP-RK3, %s)" % ("2D" if two_d or w["target"] == "heat2d_sgpu" else "3D", "WENO5 Burgers" if "burgers" in w["target"] else "heat", args.precision),
                "value": value, "unit": "Mcell-updates/s", "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
                "dtype": args.precision, "data": "synthetic", "config": config,
                "hbm_gbs": achieved * (world if nranks > 1 or w["scaling"] == "replicas" else 1),
                "roofline": roofline, "roofline_fp64": fp64, "halo": halo, "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks, "wall_s_timed_region": t_wall}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
