#!/usr/bin/env python
"""GPU baselines from the same box, multi-GPU programs: the reference's UNMODIFIED MultiGPU/*_Baseline programs
rebuilt for sm_100 against the single-node MPI stand-in (oracle/_ref/ref_*_mgpu.run, `make -C oracle refgpu`;
`mpirun -np P` -> ADV_MPI_NP=P) against this repo's drop-in drivers (ADV_NGPUS=P), same argv.

Both programs time their own loop between barriers (`Kernel time ex. data transfers`, main.c:185-321); the
reference synchronises the device before every RK launch, so at most one small kernel is still in flight when
its clock stops.  result.bin of the two runs are compared (float32, whole padded array).  One JSON object.

    python tests/baselines/ref_mgpu_baseline.py P        # P = number of GPUs / ranks
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "oracle", "_ref")
DRV = os.path.join(ROOT, "drivers")


def run(exe, args, cwd, env):
    r = subprocess.run([exe] + [str(a) for a in args], cwd=cwd, capture_output=True, text=True, env=env,
                       timeout=float(os.environ.get("REFMGPU_TIMEOUT_S", "150")))
    if r.returncode != 0:
        raise RuntimeError("%s failed (%d): %s" % (exe, r.returncode, (r.stdout + r.stderr)[-600:]))
    t = float(re.search(r"Kernel time ex\. data transfers\s*:\s*([0-9.eE+-]+)", r.stdout).group(1))
    it = float(re.search(r"iterations:\s*([0-9.eE+-]+)", r.stdout).group(1))
    return t, it


def case(name, ref_exe, our_exe, args, cells, P, out):
    only = os.environ.get("REFMGPU_ONLY")                   # e.g. REFMGPU_ONLY=Burgers3d,Diffusion3d
    if only and not any(name.startswith(x) for x in only.split(",")):
        return
    try:
        _case(name, ref_exe, our_exe, args, cells, P, out)
    except Exception as e:              # keep what finished: a hung program must not cost the whole run
        out.setdefault(name, {})["error"] = str(e)[-400:]
    print(json.dumps({name: out.get(name)}), file=sys.stderr, flush=True)


def _case(name, ref_exe, our_exe, args, cells, P, out):
    res = {"argv": " ".join(str(a) for a in args), "ranks": P}
    out[name] = res
    with tempfile.TemporaryDirectory() as d_ref, tempfile.TemporaryDirectory() as d_our:
        files = {}
        for key, exe, cwd, env in (("this_repo", os.path.join(DRV, our_exe), d_our, dict(os.environ, ADV_NGPUS=str(P))),
                                   ("reference_sm100", os.path.join(REF, ref_exe), d_ref, dict(os.environ, ADV_MPI_NP=str(P)))):
            print("[%s] %s ..." % (name, key), file=sys.stderr, flush=True)
            try:
                t, it = run(exe, args, cwd, env)
                res[key] = {"seconds": t, "iterations": it, "Mcell_updates_per_s": cells * 3 * it / t / 1e6}
                files[key] = (np.fromfile(os.path.join(cwd, "result.bin"), dtype=np.float32),
                              np.fromfile(os.path.join(cwd, "initial.bin"), dtype=np.float32))
            except Exception as e:
                res[key] = {"error": str(e)[-400:]}
        if len(files) == 2:
            (r, i0), (o, i1) = files["reference_sm100"], files["this_repo"]
            res["speedup"] = res["this_repo"]["Mcell_updates_per_s"] / res["reference_sm100"]["Mcell_updates_per_s"]
            res["result_bin_elems"] = int(r.size)
            res["result_bin_max_abs_diff"] = float(np.abs(r - o).max())
            res["result_bin_max_abs"] = float(np.abs(r).max())
            res["initial_bin_identical"] = bool(np.array_equal(i0, i1))
        # three-way check against the CPU oracle (= the reference's program-order semantics, pinned bit-exactly to
        # its kernels): tells which side a disagreement comes from
        if os.environ.get("REFMGPU_ORACLE") and name in ("Burgers3d_Baseline", "Diffusion3d_Baseline") and files:
            sys.path.insert(0, ROOT)
            from oracle import oracle as orc
            a = [float(x) for x in args]
            nx, ny, nz = int(args[-6 if name.startswith("Burgers") else -7]), int(args[-5 if name.startswith("Burgers") else -6]), int(args[-4 if name.startswith("Burgers") else -5])
            if name.startswith("Burgers"):
                want, _ = orc.burgers3d_mgpu(orc.ic_burgers3d_mgpu(nx, ny, nz, a[2], a[3], a[4]), a[0], a[1], a[2], a[3], a[4])
            else:
                want, _ = orc.heat3d(orc.ic_heat3d_mgpu(nx, ny, nz), int(args[7]), "mgpu", a[0], a[1], a[2], a[3])
            w32 = want.astype(np.float32).ravel()
            for key, (rb, _) in files.items():
                res[key]["result_bin_max_abs_diff_vs_cpu_oracle"] = float(np.abs(rb - w32).max())


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    n = int(os.environ.get("REFMGPU_N", "512"))
    n2 = int(os.environ.get("REFMGPU_N2", "4096"))
    out = {"ranks": P, "method": "each program's own 'Kernel time ex. data transfers'; same argv; reference = unmodified "
                                 "MultiGPU/*_Baseline sources, nvcc -O3 sm_100, MPI calls served by oracle/ref_harness/mpishim"}
    nz = n * P                                              # weak: n^3 per GPU
    case("Burgers3d_Baseline", "ref_Burgers3d_mgpu.run", "Burgers3d_mgpu.run",
         [float(os.environ.get("REFMGPU_TEND", "0.05")), 0.3, 2.0, 2.0, 2.0 * P, n, n, nz, 8, 8, 8], n * n * nz, P, out)
    case("Diffusion3d_Baseline", "ref_Diffusion3d_mgpu.run", "Diffusion3d_mgpu.run",
         [1.0, 2.0, 2.0, 2.0 * P, n, n, nz, int(os.environ.get("REFMGPU_ITERS", "200")), 64, 4, 1], n * n * nz, P, out)
    ny = n2 * P
    case("Burgers2d_Baseline", "ref_Burgers2d_mgpu.run", "Burgers2d_mgpu.run",
         [0.01, 0.4, 2.0, 2.0 * P, n2, ny, 32, 32], n2 * ny, P, out)
    case("Diffusion2d_Baseline", "ref_Diffusion2d_mgpu.run", "Diffusion2d_mgpu.run",
         [1.0, 2.0, 2.0 * P, n2, ny, 200, 32, 32], n2 * ny, P, out)
    print(json.dumps(out))


if __name__ == "__main__":
    sys.exit(main())
