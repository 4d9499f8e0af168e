#!/usr/bin/env python
"""GPU baselines from the same B200: the reference's own single-GPU programs rebuilt for sm_100
(oracle/_ref/ref_*.run, `make -C oracle refgpu`) against this repo's drop-in drivers, same argv.

The reference drivers time themselves with clock() and never synchronise the device before stopping the
clock, so every program -- theirs and ours -- is timed from outside as the wall-clock
DIFFERENCE between a run with 2N steps and a run with N steps (set-up, IC, transfers and file output cancel).
Where the reference writes result.bin, the two files are compared (float32).  Prints one JSON object.
"""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "oracle", "_ref")
DRV = os.path.join(ROOT, "drivers")


def wall(cmd, cwd, env=None):
    t0 = time.perf_counter()
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, env=env, timeout=900)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise RuntimeError("%s failed: %s" % (cmd[0], (r.stdout + r.stderr)[-400:]))
    return dt, r.stdout


def steps_of(out):
    for ln in out.splitlines():
        if "iterations:" in ln:
            return float(ln.split("iterations:")[1].split(",")[0])
    return None


def diff_rate(exe, args_n, args_2n, cells, steps_n, steps_2n, cwd, env=None):
    wall(exe + args_n, cwd, env)                       # warm the binary / driver
    tn = min(wall(exe + args_n, cwd, env)[0] for _ in range(3))
    t2 = min(wall(exe + args_2n, cwd, env)[0] for _ in range(3))
    d = max(t2 - tn, 1e-9)
    return {"seconds_N": tn, "seconds_2N": t2, "steps_N": steps_n, "steps_2N": steps_2n,
            "Mcell_updates_per_s": cells * 3 * (steps_2n - steps_n) / d / 1e6}


def main():
    n = int(os.environ.get("REFGPU_N", "256"))
    out = {"grid": [n, n, n], "method": "wall-clock difference between runs of 2N and N steps, best of 3 (process start-up is ~1.5 s +- 0.3 s, so the run lengths are chosen to differ by seconds), same argv for both programs"}
    env_w = dict(os.environ, ADV_WRITE="1")
    with tempfile.TemporaryDirectory() as d_ref, tempfile.TemporaryDirectory() as d_our:
        # ---- heat 3-D (C2): SingleGPU/Diffusion3d_baselineCode and _Blocking
        cells = n ** 3
        for name, ref_exe, our_exe, tail in (("heat3d_baselineCode", "ref_diffusion3d_baseline.run", "diffusion3d.run", ["32", "4", "4"]),
                                             ("heat3d_Blocking", "ref_diffusion3d_blocking.run", "diffusion3d_blocking.run", [])):
            a = lambda it: ["1.0", "10", "10", "10", str(n), str(n), str(n), str(it)] + tail
            ref = diff_rate([os.path.join(REF, ref_exe)], a(1000), a(5000), cells, 1000, 5000, d_ref)
            our = diff_rate([os.path.join(DRV, our_exe)], a(5000), a(25000), cells, 5000, 25000, d_our)
            # GPU-vs-GPU parity through result.bin (float32) on a short run
            wall([os.path.join(REF, ref_exe)] + a(20), d_ref)
            wall([os.path.join(DRV, our_exe)] + a(20), d_our)
            r = np.fromfile(os.path.join(d_ref, "result.bin"), dtype=np.float32)
            o = np.fromfile(os.path.join(d_our, "result.bin"), dtype=np.float32)
            out[name] = {"reference_sm100": ref, "this_repo": our, "speedup": our["Mcell_updates_per_s"] / ref["Mcell_updates_per_s"],
                         "result_bin_max_abs_diff_20_steps": float(np.abs(r - o).max()), "result_bin_elems": int(r.size)}
        # ---- Burgers 3-D (C3): SingleGPU/Burgers3d_WENO5 and _PitchedMem (WENO5-Z + viscous), tEnd 1 / 3
        for name, ref_exe in (("burgers3d_WENO5", "ref_burgers3d.run"), ("burgers3d_WENO5_PitchedMem", "ref_burgers3d_pitched.run")):
            a = lambda tend: [str(tend), "0.3", "2", "2", "2", str(n), str(n), str(n), "8", "8", "8"]
            _, o1 = wall([os.path.join(DRV, "burgers3d.run")] + a(1.0), d_our)
            _, o2 = wall([os.path.join(DRV, "burgers3d.run")] + a(3.0), d_our)
            s1, s2 = steps_of(o1), steps_of(o2)
            ref = diff_rate([os.path.join(REF, ref_exe)], a(1.0), a(3.0), cells, s1, s2, d_ref)
            our = diff_rate([os.path.join(DRV, "burgers3d.run")], a(1.0), a(3.0), cells, s1, s2, d_our)
            out[name] = {"reference_sm100": ref, "this_repo": our, "speedup": our["Mcell_updates_per_s"] / ref["Mcell_updates_per_s"]}
        # ---- heat 2-D (C1 on the GPU): SingleGPU/Diffusion2d, 512^2
        a = lambda it: ["1.0", "10", "10", "512", "512", str(it), "16", "16"]
        ref = diff_rate([os.path.join(REF, "ref_diffusion2d.run")], a(50000), a(250000), 512 * 512, 50000, 250000, d_ref)
        our = diff_rate([os.path.join(DRV, "diffusion2d.run")], a(50000), a(250000), 512 * 512, 50000, 250000, d_our)
        wall([os.path.join(REF, "ref_diffusion2d.run")] + a(1000), d_ref)
        wall([os.path.join(DRV, "diffusion2d.run")] + a(1000), d_our)
        r = np.fromfile(os.path.join(d_ref, "result.bin"), dtype=np.float32)
        o = np.fromfile(os.path.join(d_our, "result.bin"), dtype=np.float32)
        out["heat2d_512"] = {"reference_sm100": ref, "this_repo": our, "speedup": our["Mcell_updates_per_s"] / ref["Mcell_updates_per_s"],
                             "result_bin_max_abs_diff_1000_steps": float(np.abs(r - o).max())}
        # ---- the remaining single-GPU comparators BASELINE.json's configs name (reference side only: this repo's
        # side is the same driver as above): Diffusion3d_PitchedMem, Diffusion2d_PitchedMem, Burgers3d_WENO5_SharedMem
        # (built with its one unused texture<> declaration dropped, oracle/Makefile)
        if os.path.exists(os.path.join(REF, "ref_diffusion3d_pitched.run")):
            a = lambda it: ["1.0", "10", "10", "10", str(n), str(n), str(n), str(it), "32", "4", "4"]
            out["heat3d_PitchedMem"] = {"reference_sm100": diff_rate([os.path.join(REF, "ref_diffusion3d_pitched.run")], a(1000), a(5000), cells, 1000, 5000, d_ref)}
        if os.path.exists(os.path.join(REF, "ref_diffusion2d_pitched.run")):
            a = lambda it: ["1.0", "10", "10", "512", "512", str(it), "16", "16"]
            out["heat2d_PitchedMem_512"] = {"reference_sm100": diff_rate([os.path.join(REF, "ref_diffusion2d_pitched.run")], a(50000), a(250000), 512 * 512, 50000, 250000, d_ref)}
        if os.path.exists(os.path.join(REF, "ref_burgers3d_shared.run")):
            a = lambda tend: [str(tend), "0.3", "2", "2", "2", str(n), str(n), str(n)]
            a11 = lambda tend: a(tend) + ["8", "8", "8"]
            _, o1 = wall([os.path.join(DRV, "burgers3d.run")] + a11(1.0), d_our)
            _, o2 = wall([os.path.join(DRV, "burgers3d.run")] + a11(3.0), d_our)
            out["burgers3d_WENO5_SharedMem"] = {"reference_sm100": diff_rate([os.path.join(REF, "ref_burgers3d_shared.run")], a(1.0), a(3.0), cells, steps_of(o1), steps_of(o2), d_ref)}
    print(json.dumps(out))


if __name__ == "__main__":
    sys.exit(main())
