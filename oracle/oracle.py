"""TEST INFRASTRUCTURE ONLY -- ctypes front end of oracle/libadvdiff_oracle.so,
the CPU restatement of the reference's time-stepping path (advdiff_oracle.h).

Allowed importers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline and
--impl reference legs.  The product package never imports this module.

Arrays are numpy, C-contiguous, shaped (nz, ny, nx) / (ny, nx): x fastest, the
reference's o = i + nx*j + nx*ny*k.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libadvdiff_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path, mode=os.RTLD_LOCAL)
    return _LIB


_LIB_FMA = None


class fma_variant:
    """Context manager: inside it every routine of this module runs the SAME C restatement compiled with
    -mfma -ffp-contract=fast (oracle/Makefile, libadvdiff_oracle_fma.so).  That build is not a parity target; the
    distance between it and the default (contraction off) build is the rounding bracket of a run."""

    def __enter__(self):
        global _LIB, _LIB_FMA
        if _LIB_FMA is None:
            path = os.path.join(_HERE, "libadvdiff_oracle_fma.so")
            if not os.path.exists(path):
                build()
            _LIB_FMA = C.CDLL(path, mode=os.RTLD_LOCAL)
        self._saved = lib()
        _LIB = _LIB_FMA
        return self

    def __exit__(self, *exc):
        global _LIB
        _LIB = self._saved
        return False


def _sfx(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float64:
        return "f64", C.c_double
    if dtype == np.float32:
        return "f32", C.c_float
    raise TypeError(dtype)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def num_threads():
    return lib().orc_num_threads()


def use_all_cores():
    """OpenMP thread count := the cores this process may run on (torchrun exports OMP_NUM_THREADS=1).  Both
    oracle libraries share one libgomp, so this also covers matlab_port.c.  Returns the count in force."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    lib().orc_set_num_threads(int(n))
    return num_threads()


# ---------------------------------------------------------------- ICs
def ic_heat3d_sgpu(nx, ny, nz, L=10, W=10, H=10, dtype=np.float64):
    s, R = _sfx(dtype)
    u = np.empty((nz, ny, nx), dtype)
    f = getattr(lib(), "orc_ic_heat3d_sgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R]
    f.restype = None
    T = np.dtype(dtype).type
    # the driver parses L,W,H with atoi (SingleGPU/Diffusion3d_baselineCode/main.cpp:35-37)
    f(_ptr(u), nx, ny, nz, T(int(L)) / T(nx - 1), T(int(W)) / T(ny - 1), T(int(H)) / T(nz - 1))
    return u


def ic_heat2d_sgpu(nx, ny, L=10, W=10, dtype=np.float64):
    s, R = _sfx(dtype)
    u = np.empty((ny, nx), dtype)
    f = getattr(lib(), "orc_ic_heat2d_sgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, R, R]
    f.restype = None
    T = np.dtype(dtype).type
    f(_ptr(u), nx, ny, T(int(L)) / T(nx - 1), T(int(W)) / T(ny - 1))
    return u


def ic_burgers3d_sgpu(nx, ny, nz, L=2.0, W=2.0, H=2.0, dtype=np.float64):
    s, R = _sfx(dtype)
    u = np.empty((nz, ny, nx), dtype)
    f = getattr(lib(), "orc_ic_burgers3d_sgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R]
    f.restype = None
    T = np.dtype(dtype).type
    f(_ptr(u), nx, ny, nz, T(L) / T(nx - 1), T(W) / T(ny - 1), T(H) / T(nz - 1))
    return u


def ic_burgers3d_mgpu(nx, ny, nz, L=2.0, W=2.0, H=2.0, dtype=np.float64):
    """Padded (nz+6, ny, nx) array; dz = H/(nz-1) with the PHYSICAL nz (main.c:63)."""
    s, R = _sfx(dtype)
    u = np.empty((nz + 6, ny, nx), dtype)
    f = getattr(lib(), "orc_ic_burgers3d_mgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R]
    f.restype = None
    T = np.dtype(dtype).type
    f(_ptr(u), nx, ny, nz + 6, T(L) / T(nx - 1), T(W) / T(ny - 1), T(H) / T(nz - 1))
    return u


def ic_heat3d_mgpu(nx, ny, nz, dtype=np.float64):
    s, R = _sfx(dtype)
    u = np.empty((nz + 6, ny, nx), dtype)
    f = getattr(lib(), "orc_ic_heat3d_mgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    f.restype = None
    f(_ptr(u), nx, ny, nz + 6)
    return u


# ---------------------------------------------------------------- solvers
def burgers3d_mgpu(u0, tEnd, CFL=0.3, L=2.0, W=2.0, H=2.0, max_steps=-1):
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    NZ, ny, nx = u.shape
    f = getattr(lib(), "orc_burgers3d_mgpu_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R, R, R, C.c_int] + [C.POINTER(C.c_double)] * 3
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    rc = f(_ptr(u), nx, ny, NZ - 6, L, W, H, CFL, tEnd, max_steps, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)


def burgers3d_mgpu_cfl(u0, tEnd, CFL=0.3, L=2.0, W=2.0, H=2.0, max_steps=-1):
    """burgers3d_mgpu with dt = CFL*dx/max|u| recomputed every step (the Matlab prototype's time step)."""
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    NZ, ny, nx = u.shape
    f = getattr(lib(), "orc_burgers3d_mgpu_run_cfl_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R, R, R, C.c_int] + [C.POINTER(C.c_double)] * 3
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    rc = f(_ptr(u), nx, ny, NZ - 6, L, W, H, CFL, tEnd, max_steps, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)


def burgers3d_sgpu(u0, tEnd, CFL=0.3, L=2.0, W=2.0, H=2.0, visc_quirk=1, max_steps=-1):
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    nz, ny, nx = u.shape
    f = getattr(lib(), "orc_burgers3d_sgpu_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R, R, R, C.c_int, C.c_int] + [C.POINTER(C.c_double)] * 3
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    rc = f(_ptr(u), nx, ny, nz, L, W, H, CFL, tEnd, visc_quirk, max_steps, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)


HEAT3D_VARIANTS = {"baseline": 0, "blocking": 1, "mgpu": 2}


def heat3d(u0, iters, variant="baseline", C_=1.0, L=10, W=10, H=10):
    """variant 'mgpu': u0 is the padded (nz+6, ny, nx) array, L/W/H are doubles;
    otherwise L/W/H are integer-truncated like the single-GPU drivers' atoi."""
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    v = HEAT3D_VARIANTS[variant]
    nzs, ny, nx = u.shape
    nz = nzs - 6 if v == 2 else nzs
    if v != 2:
        L, W, H = int(L), int(W), int(H)
    f = getattr(lib(), "orc_heat3d_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, R, R, R, R, C.c_int, C.c_int,
                  C.POINTER(C.c_double), C.POINTER(C.c_int)]
    dt, it = C.c_double(), C.c_int()
    rc = f(_ptr(u), nx, ny, nz, L, W, H, C_, iters, v, C.byref(dt), C.byref(it))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value)


def heat2d(u0, iters, C_=1.0, L=10, W=10):
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    ny, nx = u.shape
    f = getattr(lib(), "orc_heat2d_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, R, R, R, C.c_int, C.POINTER(C.c_double)]
    dt = C.c_double()
    rc = f(_ptr(u), nx, ny, int(L), int(W), C_, iters, C.byref(dt))
    assert rc == 0
    return u, dict(dt=dt.value)


# ---------------------------------------------------------------- 2-D multi-GPU programs (y-slabs)
def ic_heat2d_mgpu(nx, ny, dtype=np.float64):
    """Padded (ny+6, nx) array."""
    s, R = _sfx(dtype)
    u = np.empty((ny + 6, nx), dtype)
    f = getattr(lib(), "orc_ic_heat2d_mgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int]
    f.restype = None
    f(_ptr(u), nx, ny + 6)
    return u


def ic_burgers2d_mgpu(nx, ny, L=2.0, W=2.0, dtype=np.float64):
    """Padded (ny+6, nx) array; dy = W/(ny-1) with the PHYSICAL ny (main.c:58)."""
    s, R = _sfx(dtype)
    u = np.empty((ny + 6, nx), dtype)
    f = getattr(lib(), "orc_ic_burgers2d_mgpu_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, R, R]
    f.restype = None
    T = np.dtype(dtype).type
    f(_ptr(u), nx, ny + 6, T(L) / T(nx - 1), T(W) / T(ny - 1))
    return u


def heat2d_mgpu(u0, iters, K=1.0, L=2.0, W=2.0):
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    NY, nx = u.shape
    f = getattr(lib(), "orc_heat2d_mgpu_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, R, R, R, C.c_int,
                  C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_double)]
    dt, it, t = C.c_double(), C.c_int(), C.c_double()
    rc = f(_ptr(u), nx, NY - 6, L, W, K, iters, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)


def burgers2d_mgpu(u0, tEnd, CFL=0.4, L=2.0, W=2.0, max_steps=-1):
    s, R = _sfx(u0.dtype)
    u = np.array(u0, copy=True, order="C")
    NY, nx = u.shape
    f = getattr(lib(), "orc_burgers2d_mgpu_run_" + s)
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, R, R, R, R, C.c_int] + [C.POINTER(C.c_double)] * 3
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    rc = f(_ptr(u), nx, NY - 6, L, W, CFL, tEnd, max_steps, C.byref(dt), C.byref(it), C.byref(t))
    assert rc == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)


def slab_split(h_q, n, _Nz):
    s, R = _sfx(h_q.dtype)
    NZ, ny, nx = h_q.shape
    out = np.empty((_Nz + 6, ny, nx), h_q.dtype)
    f = getattr(lib(), "orc_slab_split_" + s)
    f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 4
    f.restype = None
    f(_ptr(np.ascontiguousarray(h_q)), _ptr(out), n, nx, ny, _Nz)
    return out


def slab_merge(h_s_q, h_q, n, _Nz):
    s, R = _sfx(h_q.dtype)
    NZ, ny, nx = h_q.shape
    f = getattr(lib(), "orc_slab_merge_" + s)
    f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 4
    f.restype = None
    f(_ptr(np.ascontiguousarray(h_s_q)), _ptr(h_q), n, nx, ny, _Nz)


def save_binary(u, path):
    s, R = _sfx(u.dtype)
    f = getattr(lib(), "orc_save_binary_" + s)
    f.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p]
    assert f(_ptr(np.ascontiguousarray(u)), u.size, path.encode()) == 0


def rel_err(a, b):
    """Parity metric of BASELINE.md section 3: max|a-b| / max|b|."""
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / np.max(np.abs(b)))


# ---------------------------------------------------------------- Oracle B: C transcription of the Matlab prototypes
_MLB = None


def mlb():
    """oracle/libmatlab_port.so (matlab_port.c) -- second opinion only, see its header."""
    global _MLB
    if _MLB is None:
        path = os.path.join(_HERE, "libmatlab_port.so")
        if not os.path.exists(path):
            build()
        _MLB = C.CDLL(path, mode=os.RTLD_LOCAL)
    return _MLB


def mlb_heat3d(u0, nsteps, D=1.0, L=10.0, W=10.0, H=10.0, dt=None):
    u = np.array(u0, dtype=np.float64, copy=True, order="C")
    nz, ny, nx = u.shape
    dx, dy, dz = L / (nx - 1), W / (ny - 1), H / (nz - 1)
    if dt is None:
        dt = 1 / (2 * D * (1 / dx**2 + 1 / dy**2 + 1 / dz**2)) * 0.9          # heat3d.m:38
    f = mlb().mlb_heat3d_run
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_double] * 4 + [C.c_int]
    assert f(_ptr(u), nx, ny, nz, D / dx**2, D / dy**2, D / dz**2, dt, nsteps) == 0
    return u, dict(dt=dt)


def mlb_heat2d(u0, nsteps, D=1.0, L=10.0, W=10.0, dt=None):
    u = np.array(u0, dtype=np.float64, copy=True, order="C")
    ny, nx = u.shape
    dx, dy = L / (nx - 1), W / (ny - 1)
    if dt is None:
        dt = 1 / (2 * D * (1 / dx**2 + 1 / dy**2)) * 0.9                         # heat2d.m:36
    f = mlb().mlb_heat2d_run
    f.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 3 + [C.c_int]
    assert f(_ptr(u), nx, ny, D / dx**2, D / dy**2, dt, nsteps) == 0
    return u, dict(dt=dt)


def mlb_burgers3d(u0, tEnd, CFL=0.3, dx=None, dy=None, dz=None, dynamic_dt=False, max_steps=-1):
    u = np.array(u0, dtype=np.float64, copy=True, order="C")
    nz, ny, nx = u.shape
    f = mlb().mlb_burgers3d_run
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_double] * 5 + [C.c_int, C.c_int] + [C.POINTER(C.c_double)] * 3
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    assert f(_ptr(u), nx, ny, nz, dx, dy, dz, CFL, tEnd, int(dynamic_dt), max_steps, C.byref(dt), C.byref(it), C.byref(t)) == 0
    return u, dict(dt=dt.value, it=it.value, t=t.value)
