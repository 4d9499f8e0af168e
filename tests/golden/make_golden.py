"""Regenerates tests/golden/ref_vectors.npz from oracle/_ref, i.e. from the
REFERENCE'S OWN kernel bodies and tools replayed on the CPU (oracle/Makefile,
`make -C oracle ref`; needs /root/reference, so this only runs in the build
container).  The committed .npz is what travels: tests compare the CPU
restatement (oracle/advdiff_oracle.c) and the CUDA path against it.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

# name -> parameters; kept tiny so the fixture stays a few hundred KB
CASES = {
    # MultiGPU/Burgers3d_Baseline, WENO5-JS inviscid, P=2 emulated ranks (P>=2 is the parity semantics)
    "burgers3d_mgpu": dict(nx=20, ny=22, nz=24, L=2.0, W=2.0, H=2.0, CFL=0.3, tEnd=0.08, P=2),
    # SingleGPU/Burgers3d_WENO5, WENO5-Z + viscous term; as shipped (planes 16..18 doubled) and clean
    "burgers3d_sgpu": dict(nx=24, ny=22, nz=26, L=2.0, W=2.0, H=2.0, CFL=0.3, tEnd=0.05),
    # SingleGPU/Diffusion3d_baselineCode and _Blocking
    "heat3d_sgpu": dict(nx=24, ny=20, nz=22, L=10, W=10, H=10, C=1.0, iters=5),
    # MultiGPU/Diffusion3d_Baseline, P=2
    "heat3d_mgpu": dict(nx=16, ny=16, nz=64, L=2.0, W=2.0, H=2.0, K=1.0, iters=5, P=2),
    # SingleGPU/Diffusion2d
    "heat2d_sgpu": dict(nx=64, ny=48, L=10, W=10, C=1.0, iters=10),
    # MultiGPU/Diffusion2d_Baseline (y-slabs, disc IC), P=2
    "heat2d_mgpu": dict(nx=56, ny=48, L=2.0, W=2.0, K=1.0, iters=12, P=2),
    # MultiGPU/Burgers2d_Baseline (y-slabs, WENO5-JS inviscid), P=2
    "burgers2d_mgpu": dict(nx=60, ny=48, L=2.0, W=2.0, CFL=0.4, tEnd=0.2, P=2),
}


def main():
    assert ref.available(), "run `make -C oracle ref` first (needs /root/reference)"
    out = {}

    c = CASES["burgers3d_mgpu"]
    u0 = ref.mgpu_burgers3d_init(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"])
    u1, info = ref.mgpu_burgers3d(u0, c["P"], c["tEnd"], c["CFL"], c["L"], c["W"], c["H"])
    u1b, _ = ref.mgpu_burgers3d(u0, 4, c["tEnd"], c["CFL"], c["L"], c["W"], c["H"])
    assert np.array_equal(u1, u1b), "reference is expected to be P-invariant for P>=2"
    out["burgers3d_mgpu_ic"], out["burgers3d_mgpu_out"] = u0, u1
    out["burgers3d_mgpu_info"] = np.array([info["dt"], info["it"], info["t"]])
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "result.bin")
        ref.mgpu_save(u1, p)
        out["burgers3d_mgpu_result_bin_sha256"] = np.frombuffer(
            hashlib.sha256(open(p, "rb").read()).digest(), dtype=np.uint8)

    c = CASES["burgers3d_sgpu"]
    u0 = ref.sgpu_burgers3d_init(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"])
    ua, info = ref.sgpu_burgers3d(u0, c["tEnd"], c["CFL"], c["L"], c["W"], c["H"], laplace_mode=0)
    ub, _ = ref.sgpu_burgers3d(u0, c["tEnd"], c["CFL"], c["L"], c["W"], c["H"], laplace_mode=1)
    out["burgers3d_sgpu_ic"], out["burgers3d_sgpu_out_shipped"], out["burgers3d_sgpu_out_clean"] = u0, ua, ub
    out["burgers3d_sgpu_info"] = np.array([info["dt"], info["it"], info["t"]])

    c = CASES["heat3d_sgpu"]
    u0 = ref.sgpu_heat3d_init(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"])
    ua, dt = ref.sgpu_heat3d(u0, c["C"], c["L"], c["W"], c["H"], c["iters"], blocking=False, block=(8, 4, 4))
    ub, _ = ref.sgpu_heat3d(u0, c["C"], c["L"], c["W"], c["H"], c["iters"], blocking=True)
    out["heat3d_sgpu_ic"], out["heat3d_sgpu_out_baseline"], out["heat3d_sgpu_out_blocking"] = u0, ua, ub
    out["heat3d_sgpu_dt"] = np.array([dt])

    c = CASES["heat3d_mgpu"]
    u0 = ref.mgpu_heat3d_init(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"])
    u1, info = ref.mgpu_heat3d(u0, c["P"], c["iters"], c["K"], c["L"], c["W"], c["H"])
    u1b, _ = ref.mgpu_heat3d(u0, 1, c["iters"], c["K"], c["L"], c["W"], c["H"])
    assert np.array_equal(u1, u1b)
    out["heat3d_mgpu_ic"], out["heat3d_mgpu_out"] = u0, u1
    out["heat3d_mgpu_info"] = np.array([info["dt"], info["it"], info["t"]])

    c = CASES["heat2d_sgpu"]
    u0 = ref.sgpu_heat2d_init(c["nx"], c["ny"], c["L"], c["W"])
    u1, dt = ref.sgpu_heat2d(u0, c["C"], c["L"], c["W"], c["iters"], block=(16, 16))
    out["heat2d_sgpu_ic"], out["heat2d_sgpu_out"] = u0, u1
    out["heat2d_sgpu_dt"] = np.array([dt])

    c = CASES["heat2d_mgpu"]
    u0 = ref.mgpu_heat2d_init(c["nx"], c["ny"], c["L"], c["W"])
    u1, info = ref.mgpu_heat2d(u0, c["P"], c["iters"], c["K"], c["L"], c["W"])
    for P in (1, 4):
        assert np.array_equal(u1, ref.mgpu_heat2d(u0, P, c["iters"], c["K"], c["L"], c["W"])[0])
    out["heat2d_mgpu_ic"], out["heat2d_mgpu_out"] = u0, u1
    out["heat2d_mgpu_info"] = np.array([info["dt"], info["it"], info["t"]])

    c = CASES["burgers2d_mgpu"]
    u0 = ref.mgpu_burgers2d_init(c["nx"], c["ny"], c["L"], c["W"])
    u1, info = ref.mgpu_burgers2d(u0, c["P"], c["tEnd"], c["CFL"], c["L"], c["W"])
    for P in (1, 4):
        assert np.array_equal(u1, ref.mgpu_burgers2d(u0, P, c["tEnd"], c["CFL"], c["L"], c["W"])[0])
    out["burgers2d_mgpu_ic"], out["burgers2d_mgpu_out"] = u0, u1
    out["burgers2d_mgpu_info"] = np.array([info["dt"], info["it"], info["t"]])
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "result.bin")
        ref.mgpu2d_save(u1, p)
        out["burgers2d_mgpu_result_bin_sha256"] = np.frombuffer(
            hashlib.sha256(open(p, "rb").read()).digest(), dtype=np.uint8)

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
