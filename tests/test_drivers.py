"""The plain-C drivers (drivers/*.c over include/advdiff.h): argv / usage behaviour on the CPU, and on
the GPU the files and stdout a reference run would leave (result.bin / initial.bin: float32, headerless,
x fastest, multi-GPU files padded to Nz+6 planes)."""
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRV = os.path.join(ROOT, "drivers")


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "multigpu_advectiondiffusion_b200")])
    subprocess.check_call(["make", "-s", "-C", DRV])


@pytest.mark.parametrize("exe,usage", [
    ("burgers3d.run", "tEnd CFL L W H NX NY NZ block_x block_y block_z"),
    ("diffusion3d.run", "diffCoef L W H NX NY NZ n_iter block_x block_y block_z"),
    ("diffusion3d_blocking.run", "diffCoef L W H NX NY NZ n_iter"),
    ("diffusion2d.run", "PO CFL L W NX NY n_iter block_x block_y"),
    ("Burgers3d_mgpu.run", "tEnd CFL L W H Nx Ny NZ block_x block_y block_z"),
    ("Diffusion3d_mgpu.run", "K L W H Nx Ny NZ max_iters block_x block_y block_z"),
    ("Burgers2d_mgpu.run", "tEnd CFL L W Nx Ny block_x block_y"),
    ("Diffusion2d_mgpu.run", "tEnd K L W Nx Ny block_x block_y"),
])
def test_wrong_argc_prints_usage_and_exits_1(exe, usage):
    """Reference behaviour: printf("Usage: ...") then exit(1) (e.g. MultiGPU/Burgers3d_Baseline/main.c:50-54)."""
    _build()
    r = subprocess.run([os.path.join(DRV, exe), "1", "2"], capture_output=True, text=True)
    assert r.returncode == 1
    assert r.stdout.startswith("Usage: ") and usage in r.stdout


def test_drivers_fail_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _build()
    r = subprocess.run([os.path.join(DRV, "burgers3d.run")] + "0.1 0.3 2 2 2 32 32 32 8 8 8".split(), capture_output=True, text=True)
    assert r.returncode != 0 and "adv_create" in r.stdout


@pytest.mark.gpu
def test_burgers3d_mgpu_driver_files(tmp_path):
    _build()
    nx, ny, nz = 40, 36, 48
    args = ["0.05", "0.3", "2.0", "2.0", "2.0", str(nx), str(ny), str(nz), "8", "8", "8"]
    r = subprocess.run([os.path.join(DRV, "Burgers3d_mgpu.run")] + args, capture_output=True, text=True, cwd=tmp_path, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "=======================Burgers-3D MPI-GPU-WENO5=====================" in r.stdout
    assert "3D Grid Size                                 :  %d x %d x %d" % (nx, ny, nz + 6) in r.stdout
    g = orc.ic_burgers3d_mgpu(nx, ny, nz)
    want, info = orc.burgers3d_mgpu(g, 0.05)
    assert "iterations: %g" % info["it"] in r.stdout
    ini = np.fromfile(tmp_path / "initial.bin", dtype=np.float32)
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert ini.size == res.size == nx * ny * (nz + 6)
    assert np.array_equal(ini, g.astype(np.float32).ravel())
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    # pad planes of result.bin hold the initial values
    assert np.array_equal(res.reshape(nz + 6, ny, nx)[:3], g[:3].astype(np.float32))


@pytest.mark.gpu
def test_heat_and_sgpu_drivers(tmp_path):
    _build()
    r = subprocess.run([os.path.join(DRV, "diffusion3d.run")] + "1.0 10 10 10 40 36 32 5 32 4 4".split(),
                       capture_output=True, text=True, cwd=tmp_path, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    want, _ = orc.heat3d(orc.ic_heat3d_sgpu(40, 36, 32), 5, "baseline")
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert res.size == 40 * 36 * 32 and np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    r = subprocess.run([os.path.join(DRV, "diffusion2d.run")] + "1.0 10 10 64 48 10 16 16".split(),
                       capture_output=True, text=True, cwd=tmp_path, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    want, _ = orc.heat2d(orc.ic_heat2d_sgpu(64, 48), 10)
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert res.size == 64 * 48 and np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    env = dict(os.environ, ADV_WRITE="1")
    r = subprocess.run([os.path.join(DRV, "burgers3d.run")] + "0.04 0.3 2 2 2 32 32 32 8 8 8".split(),
                       capture_output=True, text=True, cwd=tmp_path, timeout=120, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    want, info = orc.burgers3d_sgpu(orc.ic_burgers3d_sgpu(32, 32, 32), 0.04)
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6 and "iterations: %g" % info["it"] in r.stdout


@pytest.mark.gpu
def test_2d_mgpu_driver_files(tmp_path):
    """Burgers2d.run / Diffusion2d.run drop-ins (MultiGPU/*2d_Baseline/main.c): files padded to Ny+6 rows."""
    _build()
    nx, ny = 80, 64
    r = subprocess.run([os.path.join(DRV, "Burgers2d_mgpu.run")] + ["0.1", "0.4", "2.0", "2.0", str(nx), str(ny), "32", "32"],
                       capture_output=True, text=True, cwd=tmp_path, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "===========================Burgers-2D MPI-GPU-WENO5=======================" in r.stdout
    assert "3D Grid Size                                 :  %d x %d" % (nx, ny + 6) in r.stdout
    g = orc.ic_burgers2d_mgpu(nx, ny)
    want, info = orc.burgers2d_mgpu(g, 0.1)
    assert "iterations: %g" % info["it"] in r.stdout
    ini = np.fromfile(tmp_path / "initial.bin", dtype=np.float32)
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert ini.size == res.size == nx * (ny + 6)
    assert np.array_equal(ini, g.astype(np.float32).ravel())
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    assert np.array_equal(res.reshape(ny + 6, nx)[:3], g[:3].astype(np.float32))
    r = subprocess.run([os.path.join(DRV, "Diffusion2d_mgpu.run")] + ["1.0", "2.0", "2.0", str(nx), str(ny), "30", "32", "32"],
                       capture_output=True, text=True, cwd=tmp_path, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    h = orc.ic_heat2d_mgpu(nx, ny)
    want, info = orc.heat2d_mgpu(h, 30)
    assert "iterations: %d" % info["it"] in r.stdout
    res = np.fromfile(tmp_path / "result.bin", dtype=np.float32)
    assert np.array_equal(np.fromfile(tmp_path / "initial.bin", dtype=np.float32), h.astype(np.float32).ravel())
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6


@pytest.mark.gpu
def test_mgpu_drivers_on_two_gpus_match_the_oracle(tmp_path):
    """In-process ranks (one host thread per GPU, adv_peer_connect_local): all four multi-GPU drop-ins at
    ADV_NGPUS=2 against the CPU oracle through result.bin.  Regression test for ranks ending up on
    different halo paths (one-time lookups raced between the rank threads and the heat driver stalled)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    _build()
    env = dict(os.environ, ADV_NGPUS="2")

    def run(exe, args):
        r = subprocess.run([os.path.join(DRV, exe)] + [str(a) for a in args], capture_output=True, text=True, cwd=tmp_path, timeout=60, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "Process 1 -> GPU1" in r.stdout
        return np.fromfile(tmp_path / "result.bin", dtype=np.float32)

    nx, ny, nz = 64, 40, 96         # the slab interface cuts the peak of the initial Gaussian / pulse
    res = run("Burgers3d_mgpu.run", [0.05, 0.3, 2.0, 2.0, 2.0, nx, ny, nz, 8, 8, 8])
    want, _ = orc.burgers3d_mgpu(orc.ic_burgers3d_mgpu(nx, ny, nz), 0.05)
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    res = run("Diffusion3d_mgpu.run", [1.0, 2.0, 2.0, 2.0, nx, ny, nz, 12, 64, 4, 1])
    want, _ = orc.heat3d(orc.ic_heat3d_mgpu(nx, ny, nz), 12, "mgpu", L=2.0, W=2.0, H=2.0)
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    nx, ny = 96, 128
    res = run("Burgers2d_mgpu.run", [0.1, 0.4, 2.0, 2.0, nx, ny, 32, 32])
    want, _ = orc.burgers2d_mgpu(orc.ic_burgers2d_mgpu(nx, ny), 0.1)
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6
    res = run("Diffusion2d_mgpu.run", [1.0, 2.0, 2.0, nx, ny, 40, 32, 32])
    want, _ = orc.heat2d_mgpu(orc.ic_heat2d_mgpu(nx, ny), 40)
    assert np.abs(res - want.astype(np.float32).ravel()).max() <= 1e-6


def test_summary_arithmetic_reproduces_the_recorded_run_headers(tmp_path):
    """The drivers print the reference's synthetic "Total effective GFLOPs" (updates x 8 / seconds, Tools.c:240-243,
    tools.cpp:148-151).  Feeding the recorded kernel times of two Run.m headers through drivers/common.h gives the
    recorded figures: SingleGPU/Diffusion3d_baselineCode/Run.m:3-12 (6.664205) and SingleGPU/Burgers3d_WENO5/
    Run.m:15-25 (7.036459).  The heat header's "Final Time 0.329157" is t0 + 605 dt of that driver's dt formula."""
    src = tmp_path / "g.c"
    src.write_text('#include "common.h"\nconst char* adv_last_error(void){return "";}\nint main(void){printf("%lf %lf\\n", calc_gflops_steps(5.81f, 605, 201, 199, 200),'
                   ' calc_gflops_stages(39.37f, 86, 512, 512, 512)); return 0;}\n')
    exe = tmp_path / "g"
    subprocess.check_call(["gcc", "-I", DRV, "-I", os.path.join(ROOT, "include"), "-Wno-unused-function", str(src), "-o", str(exe)])
    assert subprocess.check_output([str(exe)], text=True).split() == ["6.664205", "7.036459"]
    _, info = orc.heat3d(np.zeros((200, 199, 201)), 0, "baseline", 1.0, 10, 10, 10)
    assert "%f" % (0.1 + 605 * info["dt"]) == "0.329157"
