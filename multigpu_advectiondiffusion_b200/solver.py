"""Python mirror of the solver handle in include/advdiff.h -- the object the reference's
per-solver drivers (main.cpp / main.c) would hold.  Arrays are numpy, shaped (nz, ny, nx)
with x fastest (the reference's o = i + nx*j + nx*ny*k)."""
import ctypes as C

import numpy as np

from ._lib import Desc, F32, F64, TARGETS, check, lib

_MGPU = ("heat3d_mgpu", "burgers3d_mgpu")
_MGPU_2D = ("heat2d_mgpu", "burgers2d_mgpu")


def _np_dtype(precision):
    return np.float64 if precision == F64 else np.float32


def make_desc(target, nx, ny, nz=1, lx=2.0, ly=2.0, lz=2.0, coef=0.3, precision=F64, flags=0,
              device=0, nranks=1, rank=0, comm_id=None):
    d = Desc()
    d.target = TARGETS[target] if isinstance(target, str) else int(target)
    d.precision = precision
    d.nx, d.ny, d.nz = nx, ny, nz
    d.lx, d.ly, d.lz = lx, ly, lz
    d.coef = coef
    d.flags = flags
    d.device = device
    d.nranks, d.rank = nranks, rank
    if comm_id is not None:
        C.memmove(d.comm_id, bytes(comm_id), 128)
    return d


def reference_tEnd_for_iters(dt, iters):
    """MultiGPU/Diffusion3d_Baseline/main.c:68: tEnd = dt*max_iters feeds while (t < tEnd)."""
    return dt * iters


class Solver:
    """One rank's solver: owns three device field buffers, streams and the halo communicator."""

    def __init__(self, target, nx, ny, nz=1, **kw):
        self.target = target
        self.desc = make_desc(target, nx, ny, nz, **kw)
        self.dtype = _np_dtype(self.desc.precision)
        self._h = C.c_void_p()
        check(lib().adv_create(C.byref(self.desc), C.byref(self._h)))
        self.nx, self.ny = nx, ny
        n = lib().adv_local_elems(self._h)
        if target == "heat2d_sgpu":
            self.local_shape = (ny, nx)
        elif target in _MGPU_2D:
            self.local_shape = (n // nx, nx)        # y-slab: ny/nranks + 6 stored rows
        else:
            self.local_shape = (n // (nx * ny), ny, nx)

    # -- lifetime
    def close(self):
        if self._h:
            check(lib().adv_destroy(self._h))
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- data
    @property
    def dt(self):
        return lib().adv_dt(self._h)

    def upload(self, u):
        u = np.ascontiguousarray(u, dtype=self.dtype)
        assert u.shape == tuple(self.local_shape), (u.shape, self.local_shape)
        self._keep = u
        check(lib().adv_upload(self._h, u.ctypes.data_as(C.c_void_p)))
        check(lib().adv_sync(self._h))

    def download(self, out=None):
        if out is None:
            out = np.empty(self.local_shape, self.dtype)
        check(lib().adv_download(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def init_reference_ic(self):
        check(lib().adv_init_reference_ic(self._h))

    def init_reference_ic_device(self):
        """same closed forms evaluated on the device (throughput runs; last-bit differences from the host libm)"""
        check(lib().adv_init_reference_ic_device(self._h))

    def halo_probe(self, reps=50):
        """(ms per exchange, bytes this rank sends per exchange): the 3-plane exchange of one stage, alone"""
        ms, b = C.c_float(), C.c_double()
        check(lib().adv_halo_probe(self._h, reps, C.byref(ms), C.byref(b)))
        return ms.value, b.value

    # -- fused NVLink halo exchange (ranks = processes): exchange the 64-byte handles any way you like
    def peer_handle(self):
        h = (C.c_ubyte * 64)()
        check(lib().adv_peer_handle(self._h, C.byref(h)))
        return bytes(h)

    def peer_connect(self, left_handle, right_handle):
        check(lib().adv_peer_connect(self._h, left_handle, right_handle))

    @property
    def halo_mode(self):
        return {0: "none", 1: "nccl", 2: "peer"}[lib().adv_halo_mode(self._h)]

    # -- time stepping
    def step(self, nsteps=1):
        check(lib().adv_step(self._h, nsteps))

    def run_until(self, tEnd, max_steps=-1):
        dt, it, t = C.c_double(), C.c_double(), C.c_double()
        check(lib().adv_run_until(self._h, tEnd, max_steps, C.byref(dt), C.byref(it), C.byref(t)))
        return dict(dt=dt.value, it=it.value, t=t.value)

    def run_until_cfl(self, tEnd, max_steps=-1):
        """dt = CFL*dx/max|u| recomputed every step (the Matlab prototype's time step; Burgers targets, one rank)."""
        dt, it, t = C.c_double(), C.c_double(), C.c_double()
        check(lib().adv_run_until_cfl(self._h, tEnd, max_steps, C.byref(dt), C.byref(it), C.byref(t)))
        return dict(dt=dt.value, it=it.value, t=t.value)

    def sync(self):
        check(lib().adv_sync(self._h))

    def elapsed_ms(self):
        ms = C.c_float()
        check(lib().adv_last_elapsed_ms(self._h, C.byref(ms)))
        return ms.value

    @property
    def kernel_launches(self):
        return lib().adv_kernel_launches(self._h)

    def device_field(self):
        p, pitch = C.c_void_p(), C.c_size_t()
        check(lib().adv_device_field(self._h, C.byref(p), C.byref(pitch)))
        return p.value, pitch.value


def solve_host(target, u0, tEnd_or_iters, **kw):
    """adv_solve_host: one run of a reference binary on host buffers (upload, loop, download,
    pad planes of MGPU targets keep their input values as Merge_domains leaves them)."""
    if target == "heat2d_sgpu":
        ny, nx = u0.shape
        nz = 1
    elif target in _MGPU_2D:
        nys, nx = u0.shape
        ny, nz = nys - 6, 1
    else:
        nzs, ny, nx = u0.shape
        nz = nzs - 6 if target in _MGPU else nzs
    d = make_desc(target, nx, ny, nz, **kw)
    dtype = _np_dtype(d.precision)
    u0 = np.ascontiguousarray(u0, dtype=dtype)
    out = np.empty_like(u0)
    dt, it, t = C.c_double(), C.c_double(), C.c_double()
    check(lib().adv_solve_host(C.byref(d), u0.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p),
                               float(tEnd_or_iters), C.byref(dt), C.byref(it), C.byref(t)))
    return out, dict(dt=dt.value, it=it.value, t=t.value)
