"""TEST INFRASTRUCTURE ONLY -- ctypes access to oracle/_ref/*.so.

Those libraries are the reference's OWN kernel bodies and tools (compiled as
host C++ by oracle/Makefile from /root/reference, in place) replayed on the CPU
by the harnesses in oracle/ref_harness/.  They pin the CPU restatement
(oracle/advdiff_oracle.c) and generate tests/golden/.  Never imported by the
product package.
"""
import ctypes as C
import os
import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def available():
    return os.path.exists(os.path.join(_DIR, "libref_mgpu_burgers3d.so"))


def _load(name):
    return C.CDLL(os.path.join(_DIR, "libref_%s.so" % name), mode=os.RTLD_LOCAL)


def sgpu_heat3d_init(nx, ny, nz, L=10, W=10, H=10):
    lib = _load("sgpu_heat3d_baseline")
    u = np.empty((nz, ny, nx))
    lib.ref_init.argtypes = [C.c_uint] * 6 + [_dp]
    lib.ref_init(nx, ny, nz, int(L), int(W), int(H), u)
    return u


def sgpu_heat3d(u0, C_=1.0, L=10, W=10, H=10, iters=1, blocking=False, block=(32, 4, 4)):
    lib = _load("sgpu_heat3d_blocking" if blocking else "sgpu_heat3d_baseline")
    nz, ny, nx = u0.shape
    out = np.empty_like(u0)
    dt = C.c_double()
    lib.ref_run.argtypes = [C.c_double] + [C.c_uint] * 10 + [_dp, _dp, C.POINTER(C.c_double)]
    rc = lib.ref_run(C_, int(L), int(W), int(H), nx, ny, nz, iters, *block,
                     np.ascontiguousarray(u0), out, C.byref(dt))
    assert rc == 0
    return out, dt.value


def sgpu_heat2d_init(nx, ny, L=10, W=10):
    lib = _load("sgpu_heat2d")
    u = np.empty((ny, nx))
    lib.ref_init.argtypes = [C.c_uint] * 4 + [_dp]
    lib.ref_init(nx, ny, int(L), int(W), u)
    return u


def sgpu_heat2d(u0, C_=1.0, L=10, W=10, iters=1, block=(16, 16)):
    lib = _load("sgpu_heat2d")
    ny, nx = u0.shape
    out = np.empty_like(u0)
    dt = C.c_double()
    lib.ref_run.argtypes = [C.c_double] + [C.c_uint] * 7 + [_dp, _dp, C.POINTER(C.c_double)]
    rc = lib.ref_run(C_, int(L), int(W), nx, ny, iters, *block, np.ascontiguousarray(u0), out, C.byref(dt))
    assert rc == 0, "Diffusion2d replay needs Nx%bx==0 and Ny%by==0"
    return out, dt.value


def sgpu_burgers3d_init(nx, ny, nz, L=2.0, W=2.0, H=2.0):
    lib = _load("sgpu_burgers3d")
    u = np.empty((nz, ny, nx))
    lib.ref_init.argtypes = [C.c_uint] * 3 + [C.c_double] * 3 + [_dp]
    lib.ref_init(nx, ny, nz, L, W, H, u)
    return u


def sgpu_burgers3d(u0, tEnd, CFL=0.3, L=2.0, W=2.0, H=2.0, block=(8, 8, 8), laplace_mode=0):
    lib = _load("sgpu_burgers3d")
    nz, ny, nx = u0.shape
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    lib.ref_run.argtypes = ([C.c_double] * 5 + [C.c_uint] * 6 + [C.c_int, _dp, _dp] +
                            [C.POINTER(C.c_double)] * 3)
    rc = lib.ref_run(tEnd, CFL, L, W, H, nx, ny, nz, *block, laplace_mode,
                     np.ascontiguousarray(u0), out, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return out, dict(dt=dt.value, it=it.value, t=t.value)


def _mgpu_init(name, ic, nx, ny, nz, L, W, H):
    lib = _load(name)
    u = np.empty((nz + 6, ny, nx))
    lib.ref_init.argtypes = [C.c_int] + [C.c_uint] * 3 + [C.c_double] * 3 + [_dp]
    lib.ref_init(ic, nx, ny, nz, L, W, H, u)
    return u


def mgpu_burgers3d_init(nx, ny, nz, L=2.0, W=2.0, H=2.0, ic=3):
    return _mgpu_init("mgpu_burgers3d", ic, nx, ny, nz, L, W, H)


def mgpu_heat3d_init(nx, ny, nz, L=2.0, W=2.0, H=2.0, ic=1):
    return _mgpu_init("mgpu_heat3d", ic, nx, ny, nz, L, W, H)


def mgpu_burgers3d(u0, P, tEnd, CFL=0.3, L=2.0, W=2.0, H=2.0, block=(8, 8, 8)):
    """u0: padded (Nz+6, Ny, Nx). Returns padded result after the reference's while(t<tEnd) loop."""
    lib = _load("mgpu_burgers3d")
    NZ, ny, nx = u0.shape
    nz = NZ - 6
    assert nz % P == 0
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    lib.ref_run.argtypes = ([C.c_int] + [C.c_double] * 5 + [C.c_uint] * 6 + [_dp, _dp] +
                            [C.POINTER(C.c_double)] * 3)
    rc = lib.ref_run(P, tEnd, CFL, L, W, H, nx, ny, nz, *block,
                     np.ascontiguousarray(u0), out, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return out, dict(dt=dt.value, it=it.value, t=t.value)


def mgpu_heat3d_safe(nz, P):
    """True when LaplaceO4_async's unchecked first-plane write of every z-block
    (MultiGPU/Diffusion3d_Baseline/Kernels.cu:232-243) stays on planes that are
    owned (so it is an idempotent recompute) for every rank and window."""
    _Nz = nz // P
    _NZ = _Nz + 6
    nblk = (_NZ + 2) // 16 + (0 if (_NZ + 2) % 16 == 0 else 1)
    for grid_z in (nblk, 3):
        for bz in range(grid_z):
            for kstart in (3, 6, _Nz):
                k0 = max(kstart, 16 * bz)
                if k0 + 2 >= _NZ or k0 >= _Nz + 3:
                    return False
    return True


def mgpu_heat3d(u0, P, iters, K=1.0, L=2.0, W=2.0, H=2.0, block=(64, 4, 1)):
    lib = _load("mgpu_heat3d")
    NZ, ny, nx = u0.shape
    nz = NZ - 6
    assert nz % P == 0 and mgpu_heat3d_safe(nz, P), "size would make the reference kernel write out of bounds"
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_int(), C.c_double()
    lib.ref_run.argtypes = ([C.c_int] + [C.c_double] * 4 + [C.c_uint] * 7 + [_dp, _dp] +
                            [C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_double)])
    rc = lib.ref_run(P, K, L, W, H, nx, ny, nz, iters, *block,
                     np.ascontiguousarray(u0), out, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return out, dict(dt=dt.value, it=it.value, t=t.value)


def mgpu_save(u, path, which="mgpu_burgers3d"):
    lib = _load(which)
    NZ, ny, nx = u.shape
    lib.ref_save.argtypes = [_dp, C.c_uint, C.c_uint, C.c_uint, C.c_char_p]
    lib.ref_save(np.ascontiguousarray(u), nx, ny, NZ, path.encode())


# ---- 2-D multi-GPU programs (y-slab, padded to Ny+6 rows) -------------------------------------------------
def _mgpu2d_init(name, ic, nx, ny, L, W):
    lib = _load(name)
    u = np.empty((ny + 6, nx))
    lib.ref_init.argtypes = [C.c_int] + [C.c_uint] * 2 + [C.c_double] * 2 + [_dp]
    lib.ref_init(ic, nx, ny, L, W, u)
    return u


def mgpu_heat2d_init(nx, ny, L=2.0, W=2.0, ic=3):
    return _mgpu2d_init("mgpu_heat2d", ic, nx, ny, L, W)


def mgpu_burgers2d_init(nx, ny, L=2.0, W=2.0, ic=3):
    return _mgpu2d_init("mgpu_burgers2d", ic, nx, ny, L, W)


def mgpu_heat2d(u0, P, iters, K=1.0, L=2.0, W=2.0, block=(32, 32)):
    """u0: padded (Ny+6, Nx).  block[1] >= 4 so the one-block-row boundary launches cover their window."""
    lib = _load("mgpu_heat2d")
    NY, nx = u0.shape
    ny = NY - 6
    assert ny % P == 0 and block[1] >= 4
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_int(), C.c_double()
    lib.ref_run.argtypes = ([C.c_int] + [C.c_double] * 3 + [C.c_uint] * 5 + [_dp, _dp] +
                            [C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_double)])
    rc = lib.ref_run(P, K, L, W, nx, ny, iters, *block, np.ascontiguousarray(u0), out,
                     C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return out, dict(dt=dt.value, it=it.value, t=t.value)


def mgpu_burgers2d(u0, P, tEnd, CFL=0.4, L=2.0, W=2.0, block=(32, 32)):
    lib = _load("mgpu_burgers2d")
    NY, nx = u0.shape
    ny = NY - 6
    assert ny % P == 0
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    lib.ref_run.argtypes = ([C.c_int] + [C.c_double] * 4 + [C.c_uint] * 4 + [_dp, _dp] +
                            [C.POINTER(C.c_double)] * 3)
    rc = lib.ref_run(P, tEnd, CFL, L, W, nx, ny, *block, np.ascontiguousarray(u0), out,
                     C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return out, dict(dt=dt.value, it=it.value, t=t.value)


def mgpu2d_save(u, path, which="mgpu_burgers2d"):
    lib = _load(which)
    NY, nx = u.shape
    lib.ref_save.argtypes = [_dp, C.c_uint, C.c_uint, C.c_char_p]
    lib.ref_save(np.ascontiguousarray(u), nx, NY, path.encode())
