"""The run summaries the reference pasted at the top of its MultiGPU Run.m files record, besides timings on
unnamed hardware, the ITERATION COUNT of each run -- known answers for the drivers' time-loop arithmetic
(REAL-typed t/dt/it, clipped last step, floating-point `while (t < tEnd)`):

    MultiGPU/Burgers3d_Baseline/Run.m:4-14     400 x 400 x 406, tEnd 0.40, CFL 0.30   -> 267 x 3 RK steps
    MultiGPU/Burgers2d_Baseline/Run.m:4-14     400 x 406,       tEnd 0.40, CFL 0.40   -> 200 x 3
    MultiGPU/Diffusion3d_Baseline/Run.m:4-14   400 x 200 x 206, 100 iterations asked  -> 101 x 3 (the extra step)
    MultiGPU/Diffusion2d_Baseline/Run.m:4-14   400 x 406,       1000 iterations asked -> 1001 x 3

    SingleGPU/Burgers3d_WENO5/Run.m:15-37      512^3 and 1000 x 1000 x 200, tEnd 0.10, CFL 0.30 -> 86 x 3 and 167 x 3

The counts depend only on the spacings, so the Burgers 3-D cases use thin grids with the recorded nx."""
import numpy as np
import pytest

from oracle import oracle as orc


def test_oracle_reproduces_the_recorded_iteration_counts():
    _, i = orc.burgers3d_mgpu(orc.ic_burgers3d_mgpu(400, 12, 12, 2.0, 2.0, 4.0), 0.40, 0.30, 2.0, 2.0, 4.0)
    assert i["it"] == 267 and i["t"] == 0.40
    _, i = orc.burgers2d_mgpu(orc.ic_burgers2d_mgpu(400, 400), 0.40, 0.40)
    assert i["it"] == 200
    for nx, steps in ((512, 86), (1000, 167)):
        _, i = orc.burgers3d_sgpu(orc.ic_burgers3d_sgpu(nx, 8, 8), 0.10, 0.30)
        assert i["it"] == steps
    _, i = orc.heat2d_mgpu(orc.ic_heat2d_mgpu(400, 400), 1000)
    assert i["it"] == 1001
    _, i = orc.heat3d(orc.ic_heat3d_mgpu(400, 200, 200), 100, "mgpu", 1.0, 2.0, 2.0, 2.0)
    assert i["it"] == 101


@pytest.mark.gpu
def test_gpu_runs_of_the_recorded_configurations():
    """Same four configurations through the C ABI: iteration counts as recorded, fields within 1e-12 of the oracle
    (the full 400 x 400 x 406 Burgers run is checked for its step count on a thin grid; its oracle would take
    half an hour on the CPU)."""
    import multigpu_advectiondiffusion_b200 as adv
    u0 = orc.ic_burgers3d_mgpu(400, 20, 20, 2.0, 2.0, 4.0)
    got, gi = adv.solve_host("burgers3d_mgpu", u0, 0.40, lx=2.0, ly=2.0, lz=4.0, coef=0.30)
    want, wi = orc.burgers3d_mgpu(u0, 0.40, 0.30, 2.0, 2.0, 4.0)
    assert gi["it"] == wi["it"] == 267 and orc.rel_err(got, want) <= 1e-12
    u0 = orc.ic_burgers2d_mgpu(400, 400)
    got, gi = adv.solve_host("burgers2d_mgpu", u0, 0.40, coef=0.40)
    want, wi = orc.burgers2d_mgpu(u0, 0.40, 0.40)
    assert gi["it"] == wi["it"] == 200 and orc.rel_err(got, want) <= 1e-12
    u0 = orc.ic_heat2d_mgpu(400, 400)
    want, wi = orc.heat2d_mgpu(u0, 1000)
    got, gi = adv.solve_host("heat2d_mgpu", u0, wi["dt"] * 1000, coef=1.0)
    assert gi["it"] == wi["it"] == 1001 and orc.rel_err(got, want) <= 1e-12
    u0 = orc.ic_heat3d_mgpu(400, 200, 200)
    want, wi = orc.heat3d(u0, 100, "mgpu", 1.0, 2.0, 2.0, 2.0)
    got, gi = adv.solve_host("heat3d_mgpu", u0, wi["dt"] * 100, coef=1.0)
    assert gi["it"] == wi["it"] == 101 and orc.rel_err(got, want) <= 1e-12
