"""CPU-side checks of the drop-in boundary: the C-ABI library builds/loads without a GPU,
exports every symbol include/advdiff.h declares, and refuses to compute without a device
(no CPU fallback exists)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import multigpu_advectiondiffusion_b200 as adv
from multigpu_advectiondiffusion_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "advdiff.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(adv_[a-z0-9_]+)\s*\(", src)))


def test_header_is_plain_c():
    subprocess.check_call(["gcc", "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Werror", "-pedantic", HEADER])


def test_library_exports_every_declared_symbol():
    if not os.path.exists(adv.lib_path()):
        adv.build_library()
    L = C.CDLL(adv.lib_path())
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "libadvdiff.so lacks %s" % n
    # and the Python binding covers the same set
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_struct_layout_matches_header(tmp_path):
    """sizeof/offsetof of the ABI structs as the C compiler sees them == the ctypes mirror."""
    prog = tmp_path / "layout.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "advdiff.h"\n'
                    'int main(void){printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(adv_desc), offsetof(adv_desc, comm_id),'
                    ' offsetof(adv_desc, coef), sizeof(adv_stage_args), offsetof(adv_stage_args, dt), offsetof(adv_stage_args, pitch));return 0;}\n')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(_lib.Desc), _lib.Desc.comm_id.offset, _lib.Desc.coef.offset,
            C.sizeof(_lib.StageArgs), _lib.StageArgs.dt.offset, _lib.StageArgs.pitch.offset]
    assert got == want


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(adv.AdvError) as e:
        adv.Solver("burgers3d_mgpu", 32, 32, 32)
    assert e.value.code == 3          # ADV_ERR_NO_DEVICE
    with pytest.raises(adv.AdvError):
        adv.solve_host("heat3d_sgpu_baseline", np.zeros((16, 16, 16)), 1, lx=10, ly=10, lz=10, coef=1.0)


def test_product_does_not_touch_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import/link/exec it."""
    for top in ("multigpu_advectiondiffusion_b200", "drivers", "include", "tools"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".c", ".h", ".sh", "Makefile")):
                    txt = open(os.path.join(dirpath, f), errors="replace").read()
                    assert "oracle" not in txt.lower().replace("no cpu", ""), os.path.join(dirpath, f)


def test_descriptor_validation_needs_no_device():
    """adv_create refuses bad descriptors before it looks for a GPU (host logic of the boundary): too-small grids,
    slabs that do not divide or are thinner than 2*RADIUS, rank outside the communicator, slabs on a single-GPU
    target, non-positive C/K/CFL, unknown target / precision."""
    cases = [
        (("burgers3d_mgpu", 4, 4, 4), {}, "small"),
        (("heat3d_mgpu", 32, 32, 5), {}, "small"),
        (("burgers3d_mgpu", 32, 32, 30), dict(nranks=4, rank=0), "divisible"),
        (("burgers3d_mgpu", 32, 32, 16), dict(nranks=4, rank=0), "thinner"),
        (("burgers3d_mgpu", 32, 32, 32), dict(nranks=2, rank=2), "rank"),
        (("heat3d_sgpu_baseline", 32, 32, 32), dict(nranks=2, rank=0), "MGPU"),
        (("burgers2d_mgpu", 64, 50, 1), dict(nranks=4, rank=0), "divisible"),
        (("heat2d_mgpu", 64, 16, 1), dict(nranks=4, rank=1), "thinner"),
        (("burgers3d_mgpu", 32, 32, 32), dict(coef=0.0), "positive"),
        ((99, 32, 32, 32), {}, "target"),
        (("burgers3d_mgpu", 32, 32, 32), dict(precision=7), "precision"),
    ]
    for args, kw, word in cases:
        with pytest.raises(adv.AdvError) as e:
            adv.Solver(*args, **kw)
        assert e.value.code == 1 and word in str(e.value), (args, kw, str(e.value))


def test_null_and_state_errors_need_no_device():
    L = adv.lib()
    assert L.adv_create(None, None) == 1 and b"NULL" in L.adv_last_error()
    assert L.adv_step(None, 1) == 1 and L.adv_run_until_cfl(None, 1.0, -1, None, None, None) == 1
    assert L.adv_local_elems(None) == 0 and L.adv_dt(None) == 0.0 and L.adv_halo_mode(None) == 0
    a = _lib.StageArgs(target=adv.TARGETS["burgers3d_mgpu"], stage=4, nx=32, ny=32, nzs=38, pitch=32)
    buf = (C.c_double * 8)()
    assert L.adv_burgers3d_stage_f64(C.byref(a), buf, buf, buf, buf, None) == 1 and b"stage" in L.adv_last_error()
    a.stage = 2
    assert L.adv_burgers3d_stage_f64(C.byref(a), buf, buf, buf, buf, None) == 1 and b"alias" in L.adv_last_error()
