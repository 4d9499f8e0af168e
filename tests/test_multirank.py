"""Multi-rank paths: world_size-2 gloo run on the CPU (host-side slab logic, exchange indices,
equivalence of shipping u instead of the reference's Lu), and -- on a box with >= 2 GPUs -- the
NCCL path through the C ABI (bit-exact rank-count invariance)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mp_slab_worker.py")


def _launch(mode, nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER, mode]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)


def test_cpu_gloo_two_ranks():
    r = _launch("cpu", 2, 29731)
    assert r.returncode == 0 and "MP_OK cpu P=2" in r.stdout and "MP_OK cpu2d P=2" in r.stdout and "MP_OK cpu_heat2d P=2" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_cpu_gloo_four_ranks():
    r = _launch("cpu", 4, 29732)
    assert r.returncode == 0 and "MP_OK cpu P=4" in r.stdout and "MP_OK cpu2d P=4" in r.stdout and "MP_OK cpu_heat2d P=4" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_gpu_nccl_two_ranks():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    r = _launch("gpu", 2, 29733)
    print("".join(ln + "\n" for ln in r.stdout.splitlines() if ln.startswith("MP_OK")))      # kept in the log with pytest -s
    assert r.returncode == 0 and r.stdout.count("MP_OK gpu ") == 10 and r.stdout.count("MP_OK gpu_cfl") == 2 and r.stdout.count("halo=peer") == 6, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_gpu_four_ranks_both_halo_modes():
    """Ranks with two neighbours (P=4): NCCL and fused-peer halo paths, bit-exact against P=1."""
    import torch
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 GPUs (run under gpurun --gpus 4)")
    r = _launch("gpu", 4, 29734)
    print("".join(ln + "\n" for ln in r.stdout.splitlines() if ln.startswith("MP_OK")))
    assert r.returncode == 0 and r.stdout.count("MP_OK gpu ") == 10 and r.stdout.count("MP_OK gpu_cfl") == 2 and r.stdout.count("halo=peer") == 6, r.stdout[-3000:] + r.stderr[-3000:]
