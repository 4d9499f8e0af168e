"""The single-node MPI stand-in used to run the unmodified reference MultiGPU programs as GPU baselines
(oracle/ref_harness/mpishim; test/baseline infrastructure, never part of the product): the reference
drivers' own message pattern, on the CPU, incl. messages larger than one mailbox (32 MiB)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "ref_harness", "mpishim")


@pytest.mark.parametrize("np_,n", [(1, 1000), (2, 1000), (4, 300000), (3, 5000000)])
def test_shim_runs_the_reference_message_pattern(tmp_path, np_, n):
    exe = str(tmp_path / "check")
    subprocess.check_call(["gcc", "-O2", "-I" + SHIM, "-o", exe, os.path.join(ROOT, "tests", "mpishim_check.c"),
                           os.path.join(SHIM, "mpi_shim.c")])
    r = subprocess.run([exe, str(n)], env=dict(os.environ, ADV_MPI_NP=str(np_)), capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("SHIM_OK") == np_ and "BAD" not in r.stdout
