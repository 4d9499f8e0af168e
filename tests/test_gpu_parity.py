"""GPU parity tests proper: the CUDA path, called through the C ABI (libadvdiff.so via ctypes),
against (1) the committed outputs of the reference's own kernels (tests/golden) and (2) the CPU
oracle on the same inputs.  Tolerances are BASELINE.json's: max|u-ref|/max|ref| <= 1e-12 in
double, 1e-5 in float, over the WHOLE output array including pad planes."""
import ctypes as C

import numpy as np
import pytest

import multigpu_advectiondiffusion_b200 as adv
from multigpu_advectiondiffusion_b200 import _lib
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

TOL64 = 1e-12
TOL32 = 1e-5


def rel(a, b):
    return orc.rel_err(a, b)


# ------------------------------------------------------------------ golden (reference kernels)
def test_golden_burgers3d_mgpu(golden, golden_cases):
    c = golden_cases["burgers3d_mgpu"]
    out, info = adv.solve_host("burgers3d_mgpu", golden["burgers3d_mgpu_ic"], c["tEnd"],
                               lx=c["L"], ly=c["W"], lz=c["H"], coef=c["CFL"])
    assert rel(out, golden["burgers3d_mgpu_out"]) <= TOL64
    assert [info["dt"], info["it"], info["t"]] == list(golden["burgers3d_mgpu_info"])
    # pad planes of the merged output hold the IC bit for bit (Merge_domains)
    assert np.array_equal(out[:3], golden["burgers3d_mgpu_ic"][:3]) and np.array_equal(out[-3:], golden["burgers3d_mgpu_ic"][-3:])


def test_golden_burgers3d_sgpu(golden, golden_cases):
    c = golden_cases["burgers3d_sgpu"]
    for flags, key in ((0, "burgers3d_sgpu_out_shipped"), (adv.FLAG_VISC_ONCE, "burgers3d_sgpu_out_clean")):
        out, info = adv.solve_host("burgers3d_sgpu", golden["burgers3d_sgpu_ic"], c["tEnd"],
                                   lx=c["L"], ly=c["W"], lz=c["H"], coef=c["CFL"], flags=flags)
        assert rel(out, golden[key]) <= TOL64
    assert [info["dt"], info["it"], info["t"]] == list(golden["burgers3d_sgpu_info"])


def test_golden_heat3d(golden, golden_cases):
    c = golden_cases["heat3d_sgpu"]
    for variant in ("baseline", "blocking"):
        out, info = adv.solve_host("heat3d_sgpu_" + variant, golden["heat3d_sgpu_ic"], c["iters"],
                                   lx=c["L"], ly=c["W"], lz=c["H"], coef=c["C"])
        assert rel(out, golden["heat3d_sgpu_out_" + variant]) <= TOL64
        assert info["dt"] == golden["heat3d_sgpu_dt"][0]
    c = golden_cases["heat3d_mgpu"]
    dt = golden["heat3d_mgpu_info"][0]
    out, info = adv.solve_host("heat3d_mgpu", golden["heat3d_mgpu_ic"], dt * c["iters"],
                               lx=c["L"], ly=c["W"], lz=c["H"], coef=c["K"])
    assert rel(out, golden["heat3d_mgpu_out"]) <= TOL64
    assert info["it"] == golden["heat3d_mgpu_info"][1]


def test_golden_heat2d(golden, golden_cases):
    c = golden_cases["heat2d_sgpu"]
    out, info = adv.solve_host("heat2d_sgpu", golden["heat2d_sgpu_ic"], c["iters"], lx=c["L"], ly=c["W"], coef=c["C"])
    assert rel(out, golden["heat2d_sgpu_out"]) <= TOL64
    assert info["dt"] == golden["heat2d_sgpu_dt"][0]


# ------------------------------------------------------------------ oracle, more shapes
@pytest.mark.parametrize("shape", [(32, 32, 32), (40, 48, 56), (67, 33, 21), (129, 20, 18)])
def test_burgers3d_mgpu_vs_oracle(shape):
    nx, ny, nz = shape
    u0 = orc.ic_burgers3d_mgpu(nx, ny, nz)
    want, wi = orc.burgers3d_mgpu(u0, 0.05)
    got, gi = adv.solve_host("burgers3d_mgpu", u0, 0.05, coef=0.3)
    assert gi == wi
    assert rel(got, want) <= TOL64


@pytest.mark.parametrize("shape", [(32, 32, 32), (40, 48, 56), (70, 30, 25)])
@pytest.mark.parametrize("flags", [0, 1])
def test_burgers3d_sgpu_vs_oracle(shape, flags):
    nx, ny, nz = shape
    u0 = orc.ic_burgers3d_sgpu(nx, ny, nz)
    want, wi = orc.burgers3d_sgpu(u0, 0.04, visc_quirk=0 if flags else 1)
    got, gi = adv.solve_host("burgers3d_sgpu", u0, 0.04, coef=0.3, flags=flags)
    assert gi == wi
    assert rel(got, want) <= TOL64


@pytest.mark.parametrize("variant", ["baseline", "blocking"])
@pytest.mark.parametrize("shape", [(32, 32, 32), (40, 48, 56), (201, 19, 20)])
def test_heat3d_sgpu_vs_oracle(variant, shape):
    nx, ny, nz = shape
    u0 = orc.ic_heat3d_sgpu(nx, ny, nz)
    want, wi = orc.heat3d(u0, 10, variant)
    got, gi = adv.solve_host("heat3d_sgpu_" + variant, u0, 10, lx=10, ly=10, lz=10, coef=1.0)
    assert gi["dt"] == wi["dt"]
    assert rel(got, want) <= TOL64


@pytest.mark.parametrize("shape", [(32, 32, 32), (24, 40, 100)])
def test_heat3d_mgpu_vs_oracle(shape):
    nx, ny, nz = shape
    u0 = orc.ic_heat3d_mgpu(nx, ny, nz)
    want, wi = orc.heat3d(u0, 100, "mgpu", L=2.0, W=2.0, H=2.0)
    got, gi = adv.solve_host("heat3d_mgpu", u0, wi["dt"] * 100, coef=1.0)
    assert gi["it"] == wi["it"]
    assert rel(got, want) <= TOL64
    # frame and pad planes are exactly frozen for the 0/1 pulse (SURVEY probe P6)
    assert np.array_equal(got[:, :3], u0[:, :3]) and np.array_equal(got[:3], u0[:3])


def test_heat2d_512_full_config_vs_oracle():
    """BASELINE config C1 at full size: 512x512, 1000 steps (the oracle needs ~2 s)."""
    u0 = orc.ic_heat2d_sgpu(512, 512)
    want, wi = orc.heat2d(u0, 1000)
    got, gi = adv.solve_host("heat2d_sgpu", u0, 1000, lx=10, ly=10, coef=1.0)
    assert gi["dt"] == wi["dt"]
    assert rel(got, want) <= TOL64


def test_float_paths_vs_oracle():
    u0 = orc.ic_burgers3d_mgpu(32, 32, 32, dtype=np.float32)
    want, _ = orc.burgers3d_mgpu(u0, 0.05)
    got, _ = adv.solve_host("burgers3d_mgpu", u0, 0.05, coef=0.3, precision=adv.F32)
    assert got.dtype == np.float32 and rel(got, want) <= TOL32
    h0 = orc.ic_heat3d_sgpu(32, 32, 32, dtype=np.float32)
    want, _ = orc.heat3d(h0, 10, "baseline")
    got, _ = adv.solve_host("heat3d_sgpu_baseline", h0, 10, lx=10, ly=10, lz=10, coef=1.0, precision=adv.F32)
    assert rel(got, want) <= TOL32


def test_reference_ic_generated_by_the_library_is_bit_exact():
    for target, ic, shape in (("burgers3d_mgpu", orc.ic_burgers3d_mgpu, (20, 28, 36)),
                              ("burgers3d_sgpu", orc.ic_burgers3d_sgpu, (20, 28, 36)),
                              ("heat3d_sgpu_baseline", orc.ic_heat3d_sgpu, (20, 28, 36)),
                              ("heat3d_mgpu", orc.ic_heat3d_mgpu, (20, 28, 36))):
        nx, ny, nz = shape
        kw = dict(lx=10, ly=10, lz=10, coef=1.0) if target.startswith("heat3d_sgpu") else dict(coef=0.3)
        with adv.Solver(target, nx, ny, nz, **kw) as s:
            s.init_reference_ic()
            assert np.array_equal(s.download(), ic(nx, ny, nz)), target


# ------------------------------------------------------------------ stage-level entry points
def test_stage_entry_point_on_raw_device_pointers():
    """adv_burgers3d_stage_f64 on caller-owned device memory (torch only allocates it): one
    full step assembled from three stage calls == the oracle's step; split plane windows give
    bit-identical results to one window (what the multi-GPU schedule relies on)."""
    import torch
    nx, ny, nz = 40, 36, 32
    u0 = orc.ic_burgers3d_mgpu(nx, ny, nz)
    NZ = nz + 6
    dx, dy, dz = 2.0 / (nx - 1), 2.0 / (ny - 1), 2.0 / (nz - 1)
    dt = 0.3 * dx
    want, _ = orc.burgers3d_mgpu(u0, 1e9, max_steps=1)     # one full-dt step; pad planes restored
    dev = torch.device("cuda:0")
    A = torch.from_numpy(u0).to(dev)
    results = []
    for windows in ([(0, NZ)], [(0, 3), (3, 6), (6, 20), (20, NZ - 3), (NZ - 3, NZ)]):
        B, Cc = torch.empty_like(A), torch.empty_like(A)
        acc = torch.zeros((NZ, nx), dtype=A.dtype, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        for stage, (src, dst) in enumerate(((A, B), (B, Cc), (Cc, B)), start=1):
            for kb, ke in windows:
                a = _lib.StageArgs(target=adv.TARGETS["burgers3d_mgpu"], flags=0, stage=stage, nx=nx, ny=ny, nzs=NZ,
                                   pitch=nx, k_begin=kb, k_end=ke, own_lo=3, own_hi=NZ - 3, k_global0=0,
                                   dt=dt, dx=dx, dy=dy, dz=dz, kx=0, ky=0, kz=0)
                _lib.check(adv.lib().adv_burgers3d_stage_f64(C.byref(a), src.data_ptr(), A.data_ptr(), dst.data_ptr(),
                                                            acc.data_ptr(), st))
        torch.cuda.synchronize()
        results.append(B.cpu().numpy())
    assert np.array_equal(results[0], results[1])
    got = results[0].copy()
    got[:3], got[-3:] = u0[:3], u0[-3:]
    assert rel(got, want) <= TOL64


def test_bad_arguments_are_rejected_with_a_message():
    with pytest.raises(adv.AdvError) as e:
        adv.Solver("burgers3d_mgpu", 4, 4, 4)
    assert e.value.code == 1 and "small" in str(e.value)
    with pytest.raises(adv.AdvError):
        adv.Solver("heat3d_sgpu_baseline", 32, 32, 32, nranks=2, rank=0, lx=10, ly=10, lz=10, coef=1.0)
    with adv.Solver("heat3d_mgpu", 32, 32, 32, coef=1.0) as s:
        with pytest.raises(adv.AdvError) as e:
            s.step(1)
        assert e.value.code == 6


# ------------------------------------------------------------------ BASELINE-size properties
def test_full_size_burgers_256_properties():
    """C3/C5-sized field (256^3 + pads): properties that need no oracle run --
    a thin sub-slab of the same field agrees with the oracle to 1e-12 where the slab's own
    boundary cannot have reached (domain of dependence: 9 cells per step), constants stay
    constant, and the step is deterministic run to run."""
    n = 256
    with adv.Solver("burgers3d_mgpu", n, n, n, coef=0.3) as s:
        s.init_reference_ic()
        u0 = s.download()
        s.step(2)
        u2 = s.download()
        s.upload(u0)
        s.step(2)
        assert np.array_equal(s.download(), u2)
    assert np.isfinite(u2).all() and u2.max() <= 1.0 + 1e-9 and u2.min() >= -1e-9
    # thin slab around the centre plane through the oracle: 2 steps -> 18 planes of influence
    k0, k1 = 131 - 30, 131 + 30
    sub0 = np.ascontiguousarray(u0[k0:k1])
    dx = 2.0 / (n - 1)
    want = sub0.copy()
    import ctypes
    L = orc.lib()
    uo = np.empty_like(want); Lu = np.empty_like(want)
    L.orc_burgers3d_mgpu_step_f64.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_double] * 4
    for _ in range(2):
        L.orc_burgers3d_mgpu_step_f64(want.ctypes.data, uo.ctypes.data, Lu.ctypes.data, n, n, k1 - k0, dx, dx, dx, 0.3 * dx)
    core = slice(24, 36)
    assert rel(u2[k0:k1][core], want[core]) <= TOL64
    # constant field: every flux difference is exactly zero; only RK rounding may move it.
    # (single-GPU target: its frame is frozen.  The multi-GPU target is NOT constant-preserving --
    # the reference seeds f_old=0 at i=2 / j=2, Kernels.cu:254,327 -- and the oracle says the same)
    c = np.full((38, 64, 64), 0.37)
    out, _ = adv.solve_host("burgers3d_sgpu", c, 0.05, coef=0.3)
    assert np.abs(out - 0.37).max() < 1e-15
    out, _ = adv.solve_host("burgers3d_mgpu", c, 0.05, coef=0.3)
    want, _ = orc.burgers3d_mgpu(c, 0.05)
    assert np.abs(out - 0.37).max() > 1e-3 and rel(out, want) <= TOL64


@pytest.mark.parametrize("target,shape,kw", [
    ("burgers3d_mgpu", (40, 36, 32), dict(coef=0.3)),
    ("burgers3d_sgpu", (40, 36, 32), dict(coef=0.3)),
    ("heat3d_mgpu", (40, 36, 32), dict(coef=1.0)),
    ("heat3d_sgpu_blocking", (40, 36, 32), dict(coef=1.0, lx=10, ly=10, lz=10)),
    ("heat2d_sgpu", (64, 48, 1), dict(coef=1.0, lx=10, ly=10)),
    ("heat2d_mgpu", (64, 48, 1), dict(coef=1.0)),
    ("burgers2d_mgpu", (64, 48, 1), dict(coef=0.4)),
])
def test_graph_replay_is_bit_identical_to_plain_launches(target, shape, kw):
    """adv_step(n >= 12) on a small single-rank problem replays a captured CUDA graph of three time steps;
    n calls of adv_step(1) launch the same kernels one by one.  Same bits either way."""
    nx, ny, nz = shape
    rng = np.random.default_rng(11)
    with adv.Solver(target, nx, ny, nz, **kw) as a, adv.Solver(target, nx, ny, nz, **kw) as b:
        u0 = 0.5 * rng.random(a.local_shape) + 0.1
        a.upload(u0)
        b.upload(u0)
        a.step(16)                      # 1 plain step to reach the captured rotation or not, then graphs, then the rest
        for _ in range(16):
            b.step(1)
        ua, ub = a.download(), b.download()
        assert a.kernel_launches == b.kernel_launches == 48
    assert np.array_equal(ua, ub)
    assert np.isfinite(ua).all() and not np.array_equal(ua, u0)


def test_gpu_reproduces_the_reference_accuracy_log_3d():
    """The reference's only recorded numbers for this path (Matlab_Prototipes/DiffusionNd/TestingAccuracy.log:67-72,
    heat 3-D L1 / Linf errors against the analytic Gaussian; parameters recovered in tests/test_oracle_b.py):
    the same loop driven through adv_heat3d_stage_f64 on caller-owned device memory prints the same digits."""
    import torch
    log = {20: (2.75e-02, 5.04e-04), 40: (1.35e-02, 3.29e-04)}
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    for n, (l1, linf) in log.items():
        D, t0, L, tFinal = 1.0, 0.1, 10.0, 1.0
        dx = L / (n - 1)
        xs = -L / 2 + dx * np.arange(n)
        X, Y, Z = np.meshgrid(xs, xs, xs, indexing="ij")
        r2 = X**2 + Y**2 + Z**2
        uE = (t0 / tFinal) ** 1.5 * np.exp(-r2 / (4 * D * tFinal))
        A = torch.from_numpy(np.ascontiguousarray(np.exp(-r2 / (4 * D * t0)))).to(dev)
        B, Cc = torch.empty_like(A), torch.empty_like(A)
        k = D / (12 * dx * dx)
        dt = 1 / (2 * D * (3 / dx**2)) * 0.9
        t = t0
        while t < tFinal:
            for stage, (src, dst) in enumerate(((A, B), (B, Cc), (Cc, B)), start=1):
                a = _lib.StageArgs(target=adv.TARGETS["heat3d_sgpu_baseline"], flags=0, stage=stage, nx=n, ny=n, nzs=n, pitch=n,
                                   k_begin=0, k_end=n, own_lo=3, own_hi=n - 3, k_global0=0,
                                   dt=dt, dx=dx, dy=dx, dz=dx, kx=k, ky=k, kz=k)
                _lib.check(adv.lib().adv_heat3d_stage_f64(C.byref(a), src.data_ptr(), A.data_ptr(), dst.data_ptr(), st))
            A, B = B, A                     # the step's result is in B
            if t + dt > tFinal:
                dt = tFinal - t
            t = t + dt
        torch.cuda.synchronize()
        err = np.abs(uE - A.cpu().numpy())
        assert "%.2e" % (dx**3 * err.sum()) == "%.2e" % l1 and "%.2e" % err.max() == "%.2e" % linf, (n, dx**3 * err.sum(), err.max())


def test_dynamic_cfl_time_step_vs_oracle(monkeypatch):
    """adv_run_until_cfl: dt = CFL*dx/max|u| recomputed every step (LFWENO5FDM3d.m:71).  Default path:
    everything on the device -- max|u| folded into stage 3, dt computed by a one-thread kernel, the host only follows
    a status word; ADV_CFL_HOST_DT=1: the first version (separate reduction pass, dt computed on the host).  Same
    arithmetic, so the two give the same bits; both within 1e-12 of the oracle."""
    nx, ny, nz = 40, 36, 32
    u0 = 0.8 * orc.ic_burgers3d_mgpu(nx, ny, nz)
    tEnd = 0.15                     # 8 dynamic steps (the last one 0.0139 long, far from a borderline count), 10 fixed ones
    want, wi = orc.burgers3d_mgpu_cfl(u0, tEnd)
    _, fixed = orc.burgers3d_mgpu(u0, tEnd)
    results = {}
    for mode in ("device", "host"):
        if mode == "host":
            monkeypatch.setenv("ADV_CFL_HOST_DT", "1")
        else:
            monkeypatch.delenv("ADV_CFL_HOST_DT", raising=False)
        with adv.Solver("burgers3d_mgpu", nx, ny, nz, coef=0.3) as s:
            s.upload(u0)
            gi = s.run_until_cfl(tEnd)
            got = s.download()
            # the solver is usable afterwards: two more fixed-dt steps run from the buffer the dynamic run ended in
            s.step(2)
            after = s.download()
        got[:3], got[-3:] = u0[:3], u0[-3:]
        assert gi["it"] == wi["it"] == 8 and fixed["it"] == 10
        assert abs(gi["t"] - tEnd) < 1e-12 and abs(gi["dt"] / wi["dt"] - 1) < 1e-9
        assert rel(got, want) <= 1e-12, (mode, rel(got, want))
        results[mode] = (got, after)
    monkeypatch.delenv("ADV_CFL_HOST_DT", raising=False)
    assert np.array_equal(results["device"][0], results["host"][0])
    assert np.array_equal(results["device"][1], results["host"][1])
    # a step budget ends the run as well (max_steps), on the device path too
    with adv.Solver("burgers3d_mgpu", nx, ny, nz, coef=0.3) as s:
        s.upload(u0)
        gi = s.run_until_cfl(tEnd, max_steps=3)
        got3 = s.download()
    want3, wi3 = orc.burgers3d_mgpu_cfl(u0, tEnd, max_steps=3)
    got3[:3], got3[-3:] = u0[:3], u0[-3:]
    assert gi["it"] == wi3["it"] == 3 and rel(got3, want3) <= 1e-12
    with adv.Solver("heat3d_mgpu", nx, ny, nz, coef=1.0) as h:
        h.upload(u0)
        with pytest.raises(adv.AdvError):
            h.run_until_cfl(0.01)


@pytest.mark.parametrize("target,shape", [("burgers3d_mgpu", (40, 36, 32)), ("burgers3d_sgpu", (40, 36, 32)), ("heat3d_mgpu", (40, 36, 32)),
                                          ("heat3d_sgpu_baseline", (40, 36, 32)), ("heat2d_sgpu", (64, 48, 1)),
                                          ("burgers2d_mgpu", (72, 96, 1)), ("heat2d_mgpu", (70, 96, 1))])
def test_device_side_initial_condition_matches_the_host_generator(target, shape):
    """adv_init_reference_ic_device: the reference's closed forms evaluated by a kernel.  The device's exp()
    is not the host libm's, so Gaussians agree to a few ulp (asserted: 1e-15 of the peak value 1), and the piecewise
    constant heat ICs (square pulse, disc) exactly."""
    nx, ny, nz = shape
    L = 10.0 if target in ("heat3d_sgpu_baseline", "heat2d_sgpu") else 2.0      # the single-GPU heat drivers' domain is [-5, 5]
    with adv.Solver(target, nx, ny, nz, lx=L, ly=L, lz=L, coef=0.3 if "burgers" in target else 1.0) as s:
        s.init_reference_ic()
        host = s.download()
        s.init_reference_ic_device()
        dev = s.download()
    if target in ("heat3d_mgpu", "heat2d_mgpu"):
        assert np.array_equal(host, dev)
    else:
        assert np.abs(host - dev).max() <= 1e-15
        assert np.abs(host).max() > 0.5
