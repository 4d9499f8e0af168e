"""The allocation-free Burgers stage entry point (adv_burgers3d_stage_f64_noalloc).

Some hosts cannot allocate device memory at the point where they call the library (the device
budget is owned by the embedding application, or the loop cannot afford the device-wide
synchronisation an allocation brings).  For them the library carries its own fixed scratch array;
this test drives that entry point through the C ABI and checks every stage output against the
ordinary entry point (caller-provided scratch), for two time steps in a row so that whatever the
first step leaves in the shared array would show up in the second.  Collected last on purpose:
a run that stops at the first failure has then already run every other test.  It neither needs
nor disturbs any state of the other tests, so it also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import multigpu_advectiondiffusion_b200 as adv
from multigpu_advectiondiffusion_b200 import _lib

pytestmark = pytest.mark.gpu

TOL64 = 1e-12      # the tolerance of tests/test_gpu_parity.py (the two paths are in fact bit-identical)


def _args(stage, nx, ny, NZ, pitch, dt, d):
    return _lib.StageArgs(target=adv.TARGETS["burgers3d_mgpu"], flags=0, stage=stage, nx=nx, ny=ny, nzs=NZ, pitch=pitch,
                          k_begin=0, k_end=NZ, own_lo=3, own_hi=NZ - 3, k_global0=0,
                          dt=dt, dx=d[0], dy=d[1], dz=d[2], kx=0, ky=0, kz=0)


def test_noalloc_matches_the_allocating_entry_point_two_steps_in_a_row():
    import torch
    nx, ny, nz = 48, 40, 30
    NZ = nz + 6
    d = (2.0 / (nx - 1), 2.0 / (ny - 1), 2.0 / (nz - 1))
    dt = 0.3 * d[0]
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    z, y, x = np.meshgrid(np.linspace(-1, 1, NZ), np.linspace(-1, 1, ny), np.linspace(-1, 1, nx), indexing="ij")
    u0 = np.exp(-(x * x + y * y + z * z) / 0.1) + 0.05 * np.sin(7 * x + 3 * y - 5 * z)
    st = torch.cuda.current_stream().cuda_stream
    L = adv.lib()

    # reference path: scratch allocated by the caller (here: torch), deliberately filled with junk
    A1 = torch.from_numpy(u0).to(dev); B1 = torch.empty_like(A1); C1 = torch.empty_like(A1)
    acc = torch.full((NZ, nx), 123.0, dtype=torch.float64, device=dev)
    # path under test: same buffers layout, no scratch argument
    A2 = A1.clone(); B2 = torch.empty_like(A2); C2 = torch.empty_like(A2)

    worst = 0.0
    for step in range(2):
        for stage, (s1, d1, s2, d2) in enumerate(((A1, B1, A2, B2), (B1, C1, B2, C2), (C1, B1, C2, B2)), start=1):
            a = _args(stage, nx, ny, NZ, nx, dt, d)
            _lib.check(L.adv_burgers3d_stage_f64(C.byref(a), s1.data_ptr(), A1.data_ptr(), d1.data_ptr(), acc.data_ptr(), st))
            _lib.check(L.adv_burgers3d_stage_f64_noalloc(C.byref(a), s2.data_ptr(), A2.data_ptr(), d2.data_ptr(), st))
            torch.cuda.synchronize()
            want, got = d1.cpu().numpy(), d2.cpu().numpy()
            err = np.abs(got - want).max() / np.abs(want).max()
            worst = max(worst, err)
            assert err <= TOL64, (step, stage, err)
        A1, B1 = B1, A1
        A2, B2 = B2, A2
    assert np.isfinite(worst)
    # the field really moved (the comparison is not between two untouched buffers)
    assert np.abs(A2.cpu().numpy() - u0).max() > 1e-3


@pytest.mark.parametrize("device", [0, 1])
def test_noalloc_in_plane_windows_matches_the_oracle(device):
    """Two whole time steps through the allocation-free entry point, each stage cut into three plane windows
    (as a caller overlapping a halo exchange would), checked against the CPU oracle of the reference program --
    not only against the sibling entry point.  Stage 1 clears just the rows of the window it produces, so a
    later window must not wipe what an earlier one accumulated; the j = 2 row feeds back into genuine cells
    above 1e-12, which is what makes this comparison sensitive to it.  Device 1 (when the box has one) checks
    that each GPU gets the address of its own copy of the array."""
    import torch
    from oracle import oracle as orc
    if torch.cuda.device_count() <= device:
        pytest.skip("needs %d GPUs" % (device + 1))
    nx, ny, nz = 40, 36, 30
    NZ = nz + 6
    d = (2.0 / (nx - 1), 2.0 / (ny - 1), 2.0 / (nz - 1))
    dt = 0.3 * d[0]
    u0 = orc.ic_burgers3d_mgpu(nx, ny, nz)
    want, wi = orc.burgers3d_mgpu(u0, 1e9, max_steps=2)
    assert wi["it"] == 2 and wi["dt"] == dt
    dev = torch.device("cuda:%d" % device)
    L = adv.lib()
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        A = torch.from_numpy(u0).to(dev); B = torch.empty_like(A); Cb = torch.empty_like(A)
        windows = ((0, 9), (9, 20), (20, NZ))
        for step in range(2):
            for stage, (src, dst) in enumerate(((A, B), (B, Cb), (Cb, B)), start=1):
                for kb, ke in windows:
                    a = _args(stage, nx, ny, NZ, nx, dt, d)
                    a.k_begin, a.k_end = kb, ke
                    _lib.check(L.adv_burgers3d_stage_f64_noalloc(C.byref(a), src.data_ptr(), A.data_ptr(), dst.data_ptr(), st))
            A, B = B, A
        torch.cuda.synchronize()
        got = A.cpu().numpy()
    got[:3], got[-3:] = u0[:3], u0[-3:]         # Merge_domains leaves the pad planes of the output at their IC
    assert orc.rel_err(got, want) <= TOL64


def test_noalloc_refuses_a_stream_of_another_device():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with torch.cuda.device(1):
        s1 = torch.cuda.Stream()
    torch.cuda.set_device(0)
    u = torch.zeros((2,), dtype=torch.float64, device="cuda:0")
    a = _args(1, 16, 16, 16, 16, 1e-3, (1e-2, 1e-2, 1e-2))
    rc = adv.lib().adv_burgers3d_stage_f64_noalloc(C.byref(a), u.data_ptr(), u.data_ptr(), u.data_ptr() + 8, s1.cuda_stream)
    assert rc == 1 and "current" in adv.lib().adv_last_error().decode()


def test_noalloc_refuses_fields_larger_than_its_scratch():
    import torch
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    nx, ny = 2048, 8
    NZ = _lib.NOALLOC_MAX_ELEMS // nx + 1                 # one plane too many
    u = torch.zeros((2,), dtype=torch.float64, device=dev)   # never touched: the call must fail first
    a = _args(1, nx, ny, NZ, nx, 1e-3, (1e-2, 1e-2, 1e-2))
    rc = adv.lib().adv_burgers3d_stage_f64_noalloc(C.byref(a), u.data_ptr(), u.data_ptr(), u.data_ptr() + 8, 0)
    assert rc == 1                                           # ADV_ERR_INVALID
    msg = adv.lib().adv_last_error().decode()
    assert "scratch" in msg and str(_lib.NOALLOC_MAX_ELEMS) in msg
    # and other targets are pointed to the ordinary entry point
    a.target = adv.TARGETS["burgers3d_sgpu"]
    a.nzs = 16
    assert adv.lib().adv_burgers3d_stage_f64_noalloc(C.byref(a), u.data_ptr(), u.data_ptr(), u.data_ptr() + 8, 0) == 1
