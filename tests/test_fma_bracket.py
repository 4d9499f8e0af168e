"""What rounding alone does to the reference's recorded 2-D Burgers run (hard part 3).

oracle/advdiff_oracle.c is built twice: with floating-point contraction off (the parity target, bit-identical to
the reference's own kernels compiled for the CPU) and with -mfma -ffp-contract=fast (the same source, every
a*b+c fused -- what nvcc does to any CUDA kernel by default).  Neither build is "wrong"; the distance between them
is the rounding bracket of the run.  On MultiGPU/Burgers2d_Baseline/Run.m:4-14 (400 x 406, CFL 0.40) it stays at
1e-15 until the Gaussian steepens into a shock (t ~ 0.3) and then grows by four orders of magnitude, because
WENO weight switching at the front amplifies last-bit differences (full curve: profiles/r02_fma_bracket.json,
written by tests/baselines/fma_bracket_curve.py).  The GPU tolerance of tests/test_recorded_runs.py at t = 0.40
(1e-10) rests on the numbers pinned here; up to t = 0.20 the 1e-12 bar holds with three orders to spare.
"""
from oracle import oracle as orc

POST_SHOCK_TOL = 1e-10      # used by tests/test_recorded_runs.py at t = 0.40


def _bracket(steps):
    u0 = orc.ic_burgers2d_mgpu(400, 400)
    a, ia = orc.burgers2d_mgpu(u0, 0.40, 0.40, max_steps=steps)
    with orc.fma_variant():
        b, ib = orc.burgers2d_mgpu(u0, 0.40, 0.40, max_steps=steps)
    assert ia["it"] == ib["it"] == steps
    return orc.rel_err(b, a)


def test_the_two_builds_differ_only_by_contraction():
    # a handful of steps: the builds are not the same binary (the bracket is not trivially zero) ...
    assert 0.0 < _bracket(5) < 1e-14


def test_pre_shock_bracket_is_far_below_the_parity_bar():
    assert _bracket(100) < 1e-13            # t = 0.2005: measured 1.2e-15


def test_post_shock_bracket_exceeds_the_parity_bar_and_justifies_the_stated_tolerance():
    b = _bracket(200)                       # t = 0.40: measured 1.3e-11
    assert 1e-12 < b < POST_SHOCK_TOL / 2, b
