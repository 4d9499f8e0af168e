"""Oracle A (oracle/advdiff_oracle.c, the CUDA programs restated and pinned to the reference's kernels) against
Oracle B (oracle/matlab_port.c, an independent C transcription of the reference's Matlab prototypes -- its CPU
implementation of the same path).  The two differ in operand order and in boundary treatment
(frozen frames / pad planes vs edge-replicated ghost cells and a zeroed outer layer), so they must agree to
rounding exactly where neither boundary treatment can have reached yet.  CPU only."""
import numpy as np

from oracle import oracle as orc

TOL = 1e-13


def test_heat3d_a_vs_b_whole_domain():
    # SingleGPU/Diffusion3d_baselineCode and heat3d.m + Laplace3d.m share frame (2) and boundary values (0)
    u0 = orc.ic_heat3d_sgpu(40, 36, 32)
    a, ai = orc.heat3d(u0, 25, "baseline")
    b, _ = orc.mlb_heat3d(u0, 25, dt=ai["dt"])
    assert orc.rel_err(b, a) <= TOL
    # the prototype's own dt formula (heat3d.m:38, 1/dx^2) is the driver's (main.cpp:58, 1/dx/dx) up to an ulp
    assert abs(orc.mlb_heat3d(u0, 0)[1]["dt"] / ai["dt"] - 1) < 1e-15


def test_heat2d_a_vs_b_interior():
    # SingleGPU/Diffusion2d freezes a frame of 3, Laplace2d.m one of 2: compare inside the domain of dependence
    steps = 6
    u0 = orc.ic_heat2d_sgpu(96, 80) + 0.1 * np.random.default_rng(2).random((80, 96))
    a, ai = orc.heat2d(u0, steps)
    b, _ = orc.mlb_heat2d(u0, steps, dt=ai["dt"])
    m = 3 + 2 * 3 * steps                   # stencil radius 2, three stages per step
    assert np.abs(a - b)[m:-m, m:-m].max() / np.abs(a).max() <= TOL
    assert np.abs(a - b).max() > 1e-6       # ... and the frames really differ outside it


def test_burgers3d_a_vs_b_interior():
    # MultiGPU/Burgers3d_Baseline (pad planes, i=2/j=2 artefacts) vs LFWENO5FDM3d.m (replicated ghost cells)
    nx, ny, nz, steps = 56, 52, 50, 2
    g = orc.ic_burgers3d_mgpu(nx, ny, nz) + 0.2 * np.random.default_rng(3).random((nz + 6, ny, nx))
    a, ai = orc.burgers3d_mgpu(g, 1e9, max_steps=steps)
    b, bi = orc.mlb_burgers3d(g, 1e9, CFL=0.3, dx=2.0 / (nx - 1), dy=2.0 / (ny - 1), dz=2.0 / (nz - 1), max_steps=steps)
    assert ai == bi
    m = 3 + 3 * 3 * steps                   # stencil radius 3, three stages per step
    d = np.abs(a - b)
    assert d[m:-m, m:-m, m:-m].max() / np.abs(a).max() <= TOL
    assert d.max() > 1e-6


def test_matlab_dynamic_time_step():
    """LFWENO5FDM3d.m:71: dt = CFL*dx/max|u|.  The CUDA drivers fix max|u| = 1 (main.c:193); for the Gaussian
    pulse (max 1 at t = 0, decaying) the prototype therefore takes the same first step and longer ones after."""
    n = 40
    g = orc.ic_burgers3d_sgpu(n, n, n)
    dx = 2.0 / (n - 1)
    _, fixed = orc.mlb_burgers3d(g, 0.05, CFL=0.3, dx=dx, dy=dx, dz=dx, dynamic_dt=False)
    u, dyn = orc.mlb_burgers3d(g, 0.05, CFL=0.3, dx=dx, dy=dx, dz=dx, dynamic_dt=True)
    assert dyn["t"] == fixed["t"] == 0.05 and dyn["it"] <= fixed["it"]
    assert np.isfinite(u).all() and u.max() <= 1.0 + 1e-3 and u.min() >= -1e-3      # the pulse has amplitude 1


# ---------------------------------------------------------------------------------------------------------
# The reference's only recorded numbers for this path: Matlab_Prototipes/DiffusionNd/TestingAccuracy.log:49-72,
# the L1 / Linf error tables of TestingAccuracy.m (diffusion2dTest.m / diffusion3dTest.m against the analytic
# Gaussian).  The log does not say which parameters produced it; tEnd = 1.0, factor = 0.9 and nc = 10, 20, 40, 80
# reproduce every printed digit, which pins Oracle B (and through the tests above Oracle A's interior
# arithmetic) to an executed run of the reference's own CPU code.
# ---------------------------------------------------------------------------------------------------------
LOG_2D = {10: (2.97e-01, 2.65e-02), 20: (3.42e-02, 2.96e-03), 40: (5.14e-03, 3.59e-04), 80: (2.99e-03, 2.64e-04)}
LOG_3D = {10: (3.99e-01, 7.31e-03), 20: (2.75e-02, 5.04e-04), 40: (1.35e-02, 3.29e-04), 80: (4.70e-03, 8.33e-05)}


def _accuracy_test(n, dim, tFinal=1.0, param=0.9):
    """diffusion{2,3}dTest.m:11-66 -- including its loop order: dt is clipped AFTER the step that crosses tFinal."""
    D, t0, L = 1.0, 0.1, 10.0
    dx = L / (n - 1)
    xs = -L / 2 + dx * np.arange(n)
    grids = np.meshgrid(*([xs] * dim), indexing="ij")
    r2 = sum(g**2 for g in grids)
    u = np.exp(-r2 / (4 * D * t0))
    uE = (t0 / tFinal) ** (dim / 2) * np.exp(-r2 / (4 * D * tFinal))
    dt = 1 / (2 * D * (dim / dx**2)) * param
    t = t0
    while t < tFinal:
        if dim == 3:
            u, _ = orc.mlb_heat3d(u, 1, D=D, L=L, W=L, H=L, dt=dt)
        else:
            u, _ = orc.mlb_heat2d(u, 1, D=D, L=L, W=L, dt=dt)
        if t + dt > tFinal:
            dt = tFinal - t
        t = t + dt
    err = np.abs(uE - u)
    return dx**dim * err.sum(), err.max()


def _same_3_digits(got, logged):
    return "%.2e" % got == "%.2e" % logged


def test_oracle_b_reproduces_the_reference_accuracy_log_2d():
    for n, (l1, linf) in LOG_2D.items():
        g1, ginf = _accuracy_test(n, 2)
        assert _same_3_digits(g1, l1) and _same_3_digits(ginf, linf), (n, g1, ginf)


def test_oracle_b_reproduces_the_reference_accuracy_log_3d():
    for n, (l1, linf) in LOG_3D.items():
        g1, ginf = _accuracy_test(n, 3)
        assert _same_3_digits(g1, l1) and _same_3_digits(ginf, linf), (n, g1, ginf)


def test_oracle_a_reproduces_the_3d_accuracy_log_too():
    """Same table through the CUDA-program semantics (SingleGPU/Diffusion3d_baselineCode: same frame of 2, same
    operator) stepped with the test's loop -- what tests/test_gpu_parity.py repeats on the GPU."""
    import ctypes as C
    f = orc.lib().orc_heat3d_step_f64
    f.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [C.c_double] * 4 + [C.c_int]
    for n, (l1, linf) in LOG_3D.items():
        D, t0, L, tFinal = 1.0, 0.1, 10.0, 1.0
        dx = L / (n - 1)
        xs = -L / 2 + dx * np.arange(n)
        X, Y, Z = np.meshgrid(xs, xs, xs, indexing="ij")
        r2 = X**2 + Y**2 + Z**2
        u = np.ascontiguousarray(np.exp(-r2 / (4 * D * t0)))
        uE = (t0 / tFinal) ** 1.5 * np.exp(-r2 / (4 * D * tFinal))
        uo, Lu = np.empty_like(u), np.zeros_like(u)
        k = D / (12 * dx * dx)
        dt = 1 / (2 * D * (3 / dx**2)) * 0.9
        t = t0
        while t < tFinal:
            f(u.ctypes.data, uo.ctypes.data, Lu.ctypes.data, n, n, n, k, k, k, dt, 0)
            if t + dt > tFinal:
                dt = tFinal - t
            t = t + dt
        err = np.abs(uE - u)
        assert _same_3_digits(dx**3 * err.sum(), l1) and _same_3_digits(err.max(), linf), n


def test_dynamic_cfl_step_a_vs_b():
    """dt = CFL*dx/max|u| (LFWENO5FDM3d.m:71): Oracle A's variant (the checker of adv_run_until_cfl) against the
    prototype transcription; the maximum sits at the pulse's peak, far from either boundary treatment."""
    nx, ny, nz, steps = 56, 52, 50, 2
    g = 0.8 * orc.ic_burgers3d_mgpu(nx, ny, nz)
    a, ai = orc.burgers3d_mgpu_cfl(g, 1e9, max_steps=steps)
    b, bi = orc.mlb_burgers3d(g, 1e9, CFL=0.3, dx=2.0 / (nx - 1), dy=2.0 / (ny - 1), dz=2.0 / (nz - 1), dynamic_dt=True, max_steps=steps)
    _, fixed = orc.burgers3d_mgpu(g, 1e9, max_steps=steps)
    assert ai["it"] == bi["it"] == steps and abs(ai["t"] / bi["t"] - 1) < 1e-14
    assert ai["t"] > 1.2 * fixed["t"]            # max|u| = 0.8 < 1: longer steps than the drivers' fixed dt
    m = 3 + 9 * steps
    assert np.abs(a - b)[m:-m, m:-m, m:-m].max() / np.abs(a).max() <= TOL


def test_accuracy_log_1d_table_too():
    """The 1-D table of the same log (diffusion1dTest.m + Laplace1d.m; not on the hot path, transcribed in numpy
    here only to show that the recovered parameters explain the WHOLE log: 24 of 24 numbers)."""
    log_1d = {10: (1.81e-01, 4.97e-02), 20: (4.87e-02, 1.34e-02), 40: (9.89e-03, 2.62e-03), 80: (1.50e-03, 8.34e-04)}
    for n, (l1, linf) in log_1d.items():
        D, t0, L, tFinal = 1.0, 0.1, 10.0, 1.0
        dx = L / (n - 1)
        x = -L / 2 + dx * np.arange(n)
        Dx = D / dx**2
        u = np.exp(-x**2 / (4 * D * t0))
        uE = np.sqrt(t0 / tFinal) * np.exp(-x**2 / (4 * D * tFinal))
        dt = 1 / (2 * D * (1 / dx**2)) * 0.9
        t = t0

        def lap(w):
            d = np.zeros_like(w)
            d[2:n - 2] = Dx / 12 * (-w[4:n] + 16 * w[3:n - 1] - 30 * w[2:n - 2] + 16 * w[1:n - 3] - w[0:n - 4])
            return d
        while t < tFinal:
            uo = u.copy()
            u = uo + dt * lap(u)
            u = 0.75 * uo + 0.25 * (u + dt * lap(u))
            u = (uo + 2 * (u + dt * lap(u))) / 3
            u[0] = u[-1] = 0
            if t + dt > tFinal:
                dt = tFinal - t
            t = t + dt
        err = np.abs(uE - u)
        assert _same_3_digits(dx * err.sum(), l1) and _same_3_digits(err.max(), linf), (n, dx * err.sum(), err.max())
