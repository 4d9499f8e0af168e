#!/usr/bin/env python
"""GPU baselines from the same box, multi-GPU programs: the reference's UNMODIFIED MultiGPU/*_Baseline programs
rebuilt for sm_100 against the single-node MPI stand-in (oracle/_ref/ref_*_mgpu.run, `make -C oracle refgpu`;
`mpirun -np P` -> ADV_MPI_NP=P) against this repo's drop-in drivers (ADV_NGPUS=P), same argv.

Both programs time their own loop between barriers (`Kernel time ex. data transfers`, main.c:185-321); the
reference synchronises the device before every RK launch, so at most one small kernel is still in flight when
its clock stops.  result.bin of the two runs are compared (float32, whole padded array).  One JSON object.

    python tools/ref_mgpu_baseline.py P        # P = number of GPUs / ranks
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
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
    try:
        _case(name, ref_exe, our_exe, args, cells, P, out)
    except Exception as e:              # keep what finished: a hung program must not cost the whole run
        out[name] = {"argv": " ".join(str(a) for a in args), "ranks": P, "error": str(e)[-400:]}
    print(json.dumps({name: out[name]}), file=sys.stderr, flush=True)


def _case(name, ref_exe, our_exe, args, cells, P, out):
    with tempfile.TemporaryDirectory() as d_ref, tempfile.TemporaryDirectory() as d_our:
        tr, itr = run(os.path.join(REF, ref_exe), args, d_ref, dict(os.environ, ADV_MPI_NP=str(P)))
        to, ito = run(os.path.join(DRV, our_exe), args, d_our, dict(os.environ, ADV_NGPUS=str(P)))
        r = np.fromfile(os.path.join(d_ref, "result.bin"), dtype=np.float32)
        o = np.fromfile(os.path.join(d_our, "result.bin"), dtype=np.float32)
        i0 = np.fromfile(os.path.join(d_ref, "initial.bin"), dtype=np.float32)
        i1 = np.fromfile(os.path.join(d_our, "initial.bin"), dtype=np.float32)
    out[name] = {"argv": " ".join(str(a) for a in args), "ranks": P, "iterations": [itr, ito],
                 "reference_sm100": {"seconds": tr, "Mcell_updates_per_s": cells * 3 * itr / tr / 1e6},
                 "this_repo": {"seconds": to, "Mcell_updates_per_s": cells * 3 * ito / to / 1e6},
                 "speedup": tr / to * (ito / itr),
                 "result_bin_elems": int(r.size), "result_bin_max_abs_diff": float(np.abs(r - o).max()),
                 "result_bin_max_abs": float(np.abs(r).max()), "initial_bin_identical": bool(np.array_equal(i0, i1))}


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    n = int(os.environ.get("REFMGPU_N", "512"))
    n2 = int(os.environ.get("REFMGPU_N2", "4096"))
    out = {"ranks": P, "method": "each program's own 'Kernel time ex. data transfers'; same argv; reference = unmodified "
                                 "MultiGPU/*_Baseline sources, nvcc -O3 sm_100, MPI calls served by oracle/ref_harness/mpishim"}
    nz = n * P                                              # weak: n^3 per GPU
    case("Burgers3d_Baseline", "ref_Burgers3d_mgpu.run", "Burgers3d_mgpu.run",
         [0.012, 0.3, 2.0, 2.0, 2.0 * P, n, n, nz, 8, 8, 8], n * n * nz, P, out)
    case("Diffusion3d_Baseline", "ref_Diffusion3d_mgpu.run", "Diffusion3d_mgpu.run",
         [1.0, 2.0, 2.0, 2.0 * P, n, n, nz, 30, 64, 4, 1], n * n * nz, P, out)
    ny = n2 * P
    case("Burgers2d_Baseline", "ref_Burgers2d_mgpu.run", "Burgers2d_mgpu.run",
         [0.01, 0.4, 2.0, 2.0 * P, n2, ny, 32, 32], n2 * ny, P, out)
    case("Diffusion2d_Baseline", "ref_Diffusion2d_mgpu.run", "Diffusion2d_mgpu.run",
         [1.0, 2.0, 2.0 * P, n2, ny, 200, 32, 32], n2 * ny, P, out)
    print(json.dumps(out))


if __name__ == "__main__":
    sys.exit(main())
