"""Pins the CPU restatement (oracle/advdiff_oracle.c) to the reference.

(1) against tests/golden/ref_vectors.npz -- outputs of the reference's own
    kernel bodies replayed on the CPU (tests/golden/make_golden.py); runs
    everywhere, bit-exact (both sides are IEEE double without FMA fusion);
(2) live against oracle/_ref on more sizes and rank counts, only where that
    directory was built (it needs /root/reference).
The reference itself ships no golden vectors for this path.
"""
import hashlib
import os

import numpy as np
import pytest

from oracle import oracle as orc
from oracle import ref


def test_ic_bit_exact(golden, golden_cases):
    c = golden_cases["burgers3d_mgpu"]
    assert np.array_equal(orc.ic_burgers3d_mgpu(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"]), golden["burgers3d_mgpu_ic"])
    c = golden_cases["burgers3d_sgpu"]
    assert np.array_equal(orc.ic_burgers3d_sgpu(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"]), golden["burgers3d_sgpu_ic"])
    c = golden_cases["heat3d_sgpu"]
    assert np.array_equal(orc.ic_heat3d_sgpu(c["nx"], c["ny"], c["nz"], c["L"], c["W"], c["H"]), golden["heat3d_sgpu_ic"])
    c = golden_cases["heat3d_mgpu"]
    assert np.array_equal(orc.ic_heat3d_mgpu(c["nx"], c["ny"], c["nz"]), golden["heat3d_mgpu_ic"])
    c = golden_cases["heat2d_sgpu"]
    assert np.array_equal(orc.ic_heat2d_sgpu(c["nx"], c["ny"], c["L"], c["W"]), golden["heat2d_sgpu_ic"])


def test_burgers3d_mgpu_golden(golden, golden_cases, tmp_path):
    c = golden_cases["burgers3d_mgpu"]
    out, info = orc.burgers3d_mgpu(golden["burgers3d_mgpu_ic"], c["tEnd"], c["CFL"], c["L"], c["W"], c["H"])
    assert np.array_equal(out, golden["burgers3d_mgpu_out"])
    assert np.array_equal(np.array([info["dt"], info["it"], info["t"]]), golden["burgers3d_mgpu_info"])
    # result.bin: float32, headerless, x fastest, padded to nz+6 planes (Tools.c:91-119)
    p = str(tmp_path / "result.bin")
    orc.save_binary(out, p)
    raw = open(p, "rb").read()
    assert len(raw) == 4 * out.size
    assert hashlib.sha256(raw).digest() == golden["burgers3d_mgpu_result_bin_sha256"].tobytes()
    assert raw == out.astype(np.float32).tobytes()


def test_burgers3d_sgpu_golden(golden, golden_cases):
    c = golden_cases["burgers3d_sgpu"]
    for quirk, key in ((1, "burgers3d_sgpu_out_shipped"), (0, "burgers3d_sgpu_out_clean")):
        out, info = orc.burgers3d_sgpu(golden["burgers3d_sgpu_ic"], c["tEnd"], c["CFL"], c["L"], c["W"], c["H"], visc_quirk=quirk)
        assert np.array_equal(out, golden[key])
    assert np.array_equal(np.array([info["dt"], info["it"], info["t"]]), golden["burgers3d_sgpu_info"])
    # the shipped double-count really is visible and confined to the neighbourhood of planes 16..18
    d = np.abs(golden["burgers3d_sgpu_out_shipped"] - golden["burgers3d_sgpu_out_clean"]).max(axis=(1, 2))
    assert d[16:19].max() > 0 and d[16:19].max() == d.max()


def test_heat3d_golden(golden, golden_cases):
    c = golden_cases["heat3d_sgpu"]
    for variant in ("baseline", "blocking"):
        out, info = orc.heat3d(golden["heat3d_sgpu_ic"], c["iters"], variant, c["C"], c["L"], c["W"], c["H"])
        assert np.array_equal(out, golden["heat3d_sgpu_out_" + variant])
        assert info["dt"] == golden["heat3d_sgpu_dt"][0]
    c = golden_cases["heat3d_mgpu"]
    out, info = orc.heat3d(golden["heat3d_mgpu_ic"], c["iters"], "mgpu", c["K"], c["L"], c["W"], c["H"])
    assert np.array_equal(out, golden["heat3d_mgpu_out"])
    assert info["dt"] == golden["heat3d_mgpu_info"][0] and info["it"] == golden["heat3d_mgpu_info"][1]


def test_heat2d_golden(golden, golden_cases):
    c = golden_cases["heat2d_sgpu"]
    out, info = orc.heat2d(golden["heat2d_sgpu_ic"], c["iters"], c["C"], c["L"], c["W"])
    assert np.array_equal(out, golden["heat2d_sgpu_out"])
    assert info["dt"] == golden["heat2d_sgpu_dt"][0]


def test_2d_mgpu_golden(golden, golden_cases, tmp_path):
    """MultiGPU/Diffusion2d_Baseline and MultiGPU/Burgers2d_Baseline (y-slabs)."""
    c = golden_cases["heat2d_mgpu"]
    assert np.array_equal(orc.ic_heat2d_mgpu(c["nx"], c["ny"]), golden["heat2d_mgpu_ic"])
    out, info = orc.heat2d_mgpu(golden["heat2d_mgpu_ic"], c["iters"], c["K"], c["L"], c["W"])
    assert np.array_equal(out, golden["heat2d_mgpu_out"])
    assert np.array_equal(np.array([info["dt"], info["it"], info["t"]]), golden["heat2d_mgpu_info"])
    c = golden_cases["burgers2d_mgpu"]
    assert np.array_equal(orc.ic_burgers2d_mgpu(c["nx"], c["ny"], c["L"], c["W"]), golden["burgers2d_mgpu_ic"])
    out, info = orc.burgers2d_mgpu(golden["burgers2d_mgpu_ic"], c["tEnd"], c["CFL"], c["L"], c["W"])
    assert np.array_equal(out, golden["burgers2d_mgpu_out"])
    assert np.array_equal(np.array([info["dt"], info["it"], info["t"]]), golden["burgers2d_mgpu_info"])
    p = str(tmp_path / "result.bin")
    orc.save_binary(out, p)
    assert hashlib.sha256(open(p, "rb").read()).digest() == golden["burgers2d_mgpu_result_bin_sha256"].tobytes()


def test_heat_analytic_soft_pin():
    """The reference's only numeric known answer for heat: the analytic Gaussian
    u=(t0/t)^{3/2} exp(-r^2/4t) used by every heat driver (heat3d.m:36,
    acoustics3d.h:29).  Loose by nature (truncation error + zeroed boundary)."""
    n = 41
    u0 = orc.ic_heat3d_sgpu(n, n, n)
    iters = 20
    u, info = orc.heat3d(u0, iters, "baseline")
    t = 0.1 + info["dt"] * iters
    x = -5 + np.arange(n) * (10.0 / (n - 1))
    X, Y, Z = np.meshgrid(x, x, x, indexing="ij")
    exact = (0.1 / t) ** 1.5 * np.exp(-(X**2 + Y**2 + Z**2) / (4 * t))
    assert np.abs(u - exact.transpose(2, 1, 0)).max() < 2e-2


def test_float_variant_close_to_double():
    """USE_FLOAT build of the restatement stays within the float tolerance (1e-5)."""
    u0 = orc.ic_burgers3d_mgpu(20, 20, 18)
    d, _ = orc.burgers3d_mgpu(u0, 0.05)
    f, _ = orc.burgers3d_mgpu(u0.astype(np.float32), 0.05)
    assert orc.rel_err(f, d) < 1e-5


def test_slab_split_merge_roundtrip():
    rng = np.random.default_rng(0)
    nz, P = 24, 4
    g = rng.standard_normal((nz + 6, 5, 7))
    out = np.zeros_like(g)
    for n in range(P):
        s = orc.slab_split(g, n, nz // P)
        assert np.array_equal(s, g[n * (nz // P): n * (nz // P) + nz // P + 6])
        orc.slab_merge(s, out, n, nz // P)
    assert np.array_equal(out[3:nz + 3], g[3:nz + 3]) and not out[:3].any() and not out[nz + 3:].any()


needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("shape,P,tEnd", [((24, 24, 24), 2, 0.06), ((20, 28, 36), 3, 0.05), ((28, 20, 32), 4, 0.04)])
def test_live_burgers3d_mgpu(shape, P, tEnd):
    nx, ny, nz = shape
    u0 = ref.mgpu_burgers3d_init(nx, ny, nz)
    assert np.array_equal(u0, orc.ic_burgers3d_mgpu(nx, ny, nz))
    r, ri = ref.mgpu_burgers3d(u0, P, tEnd)
    o, oi = orc.burgers3d_mgpu(u0, tEnd)
    assert np.array_equal(r, o) and ri == oi


@needs_ref
def test_live_reference_np1_is_not_the_parity_target():
    """the reference at np=1 accumulates the j=2 artefact twice."""
    u0 = ref.mgpu_burgers3d_init(24, 24, 24)
    r1, _ = ref.mgpu_burgers3d(u0, 1, 0.06)
    r2, _ = ref.mgpu_burgers3d(u0, 2, 0.06)
    assert not np.array_equal(r1, r2)


@needs_ref
@pytest.mark.parametrize("shape", [(40, 40, 40), (28, 28, 35)])
def test_live_burgers3d_sgpu(shape):
    # Nx == Ny only: the reference's x/y sweep launches swap the grid extents
    # (SingleGPU/Burgers3d_WENO5/main.cpp:99-105) and do not cover the domain otherwise.
    nx, ny, nz = shape
    u0 = ref.sgpu_burgers3d_init(nx, ny, nz)
    for mode in (0, 1):
        r, ri = ref.sgpu_burgers3d(u0, 0.04, block=(4, 4, 4), laplace_mode=mode)
        o, oi = orc.burgers3d_sgpu(u0, 0.04, visc_quirk=1 - mode)
        assert np.array_equal(r, o) and ri == oi


@needs_ref
@pytest.mark.parametrize("shape", [(32, 36, 40), (24, 24, 22)])
def test_live_heat3d_sgpu(shape):
    nx, ny, nz = shape
    u0 = ref.sgpu_heat3d_init(nx, ny, nz)
    for blocking, variant in ((False, "baseline"), (True, "blocking")):
        r, dt = ref.sgpu_heat3d(u0, iters=4, blocking=blocking)
        o, oi = orc.heat3d(u0, 4, variant)
        assert np.array_equal(r, o) and dt == oi["dt"]


@needs_ref
@pytest.mark.parametrize("P", [1, 2, 4])
def test_live_heat3d_mgpu(P):
    u0 = ref.mgpu_heat3d_init(24, 20, 128)
    r, ri = ref.mgpu_heat3d(u0, P, 4)
    o, oi = orc.heat3d(u0, 4, "mgpu", L=2.0, W=2.0, H=2.0)
    assert np.array_equal(r, o) and ri["it"] == oi["it"]


@needs_ref
def test_live_heat2d():
    u0 = ref.sgpu_heat2d_init(96, 64)
    r, dt = ref.sgpu_heat2d(u0, iters=12, block=(16, 16))
    o, oi = orc.heat2d(u0, 12)
    assert np.array_equal(r, o) and dt == oi["dt"]


@needs_ref
@pytest.mark.parametrize("shape,P,iters", [((48, 64), 1, 9), ((37, 40), 2, 7), ((80, 96), 4, 5)])
def test_live_heat2d_mgpu(shape, P, iters):
    nx, ny = shape
    u0 = ref.mgpu_heat2d_init(nx, ny)
    assert np.array_equal(u0, orc.ic_heat2d_mgpu(nx, ny))
    u0 = u0 + np.random.default_rng(3).random(u0.shape)      # exercise every row/column of the footprint
    r, ri = ref.mgpu_heat2d(u0, P, iters)
    o, oi = orc.heat2d_mgpu(u0, iters)
    assert np.array_equal(r, o) and ri == oi


@needs_ref
@pytest.mark.parametrize("shape,P,tEnd", [((48, 64), 1, 0.1), ((37, 40), 2, 0.12), ((80, 96), 4, 0.05)])
def test_live_burgers2d_mgpu(shape, P, tEnd):
    nx, ny = shape
    u0 = ref.mgpu_burgers2d_init(nx, ny)
    assert np.array_equal(u0, orc.ic_burgers2d_mgpu(nx, ny))
    for ic in (u0, np.random.default_rng(4).random(u0.shape) - 0.3):
        r, ri = ref.mgpu_burgers2d(ic, P, tEnd)
        o, oi = orc.burgers2d_mgpu(ic, tEnd)
        assert np.array_equal(r, o) and ri == oi
