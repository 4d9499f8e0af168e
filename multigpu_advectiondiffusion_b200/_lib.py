"""ctypes binding of libadvdiff.so (include/advdiff.h).  Loads the in-tree library and fails
loudly when it is missing: there is no fallback path."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBNAME = "libadvdiff.so"

TARGETS = {
    "heat2d_sgpu": 0,
    "heat3d_sgpu_baseline": 1,
    "heat3d_sgpu_blocking": 2,
    "heat3d_mgpu": 3,
    "burgers3d_sgpu": 4,
    "burgers3d_mgpu": 5,
    "heat2d_mgpu": 6,
    "burgers2d_mgpu": 7,
}
F64, F32 = 0, 1
FLAG_VISC_ONCE = 0x1
NOALLOC_MAX_ELEMS = 1179648

ERR_NAMES = {1: "INVALID", 2: "CUDA", 3: "NO_DEVICE", 4: "COMM", 5: "NOMEM", 6: "STATE"}


class AdvError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("advdiff error %d (%s): %s" % (code, ERR_NAMES.get(code, "?"), msg))
        self.code = code


class Desc(C.Structure):
    _fields_ = [("target", C.c_int), ("precision", C.c_int),
                ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int),
                ("lx", C.c_double), ("ly", C.c_double), ("lz", C.c_double),
                ("coef", C.c_double), ("flags", C.c_int), ("device", C.c_int),
                ("nranks", C.c_int), ("rank", C.c_int),
                ("comm_id", C.c_ubyte * 128)]


class StageArgs(C.Structure):
    _fields_ = [("target", C.c_int), ("flags", C.c_int), ("stage", C.c_int),
                ("nx", C.c_int), ("ny", C.c_int), ("nzs", C.c_int),
                ("pitch", C.c_size_t),
                ("k_begin", C.c_int), ("k_end", C.c_int),
                ("own_lo", C.c_int), ("own_hi", C.c_int), ("k_global0", C.c_int),
                ("dt", C.c_double),
                ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
                ("kx", C.c_double), ("ky", C.c_double), ("kz", C.c_double)]


# every symbol include/advdiff.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "adv_last_error": (C.c_char_p, []),
    "adv_version": (C.c_int, []),
    "adv_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "adv_comm_unique_id": (C.c_int, [C.POINTER(C.c_ubyte * 128)]),
    "adv_create": (C.c_int, [C.POINTER(Desc), C.POINTER(_P)]),
    "adv_destroy": (C.c_int, [_P]),
    "adv_local_elems": (C.c_size_t, [_P]),
    "adv_dt": (C.c_double, [_P]),
    "adv_upload": (C.c_int, [_P, _P]),
    "adv_download": (C.c_int, [_P, _P]),
    "adv_init_reference_ic": (C.c_int, [_P]),
    "adv_init_reference_ic_device": (C.c_int, [_P]),
    "adv_halo_probe": (C.c_int, [_P, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_double)]),
    "adv_step": (C.c_int, [_P, C.c_int]),
    "adv_run_until": (C.c_int, [_P, C.c_double, C.c_int] + [C.POINTER(C.c_double)] * 3),
    "adv_run_until_cfl": (C.c_int, [_P, C.c_double, C.c_int] + [C.POINTER(C.c_double)] * 3),
    "adv_sync": (C.c_int, [_P]),
    "adv_last_elapsed_ms": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "adv_kernel_launches": (C.c_longlong, [_P]),
    "adv_device_field": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "adv_peer_handle": (C.c_int, [_P, C.POINTER(C.c_ubyte * 64)]),
    "adv_peer_connect": (C.c_int, [_P, C.c_char_p, C.c_char_p]),
    "adv_peer_connect_local": (C.c_int, [_P, _P, _P]),
    "adv_halo_mode": (C.c_int, [_P]),
    "adv_solve_host": (C.c_int, [C.POINTER(Desc), _P, _P, C.c_double] + [C.POINTER(C.c_double)] * 3),
    "adv_heat3d_stage_f64": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_heat3d_stage_f32": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_heat2d_stage_f64": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_heat2d_stage_f32": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_burgers2d_stage_f64": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_burgers2d_stage_f32": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
    "adv_burgers3d_stage_f64": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P, _P]),
    "adv_burgers3d_stage_f32": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P, _P]),
    "adv_burgers3d_stage_f64_noalloc": (C.c_int, [C.POINTER(StageArgs), _P, _P, _P, _P]),
}

_LIB = None


def lib_path():
    return os.path.join(_HERE, _LIBNAME)


def build_library(verbose=False):
    """Compile csrc/ for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", _HERE] + ([] if verbose else ["-s"]))
    return lib_path()


def lib():
    """The loaded C-ABI library.  Raises if it has not been built -- no fallback."""
    global _LIB
    if _LIB is None:
        path = lib_path()
        if not os.path.exists(path):
            raise AdvError(3, "%s is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "or `make -C multigpu_advectiondiffusion_b200` (there is no CPU fallback)" % path)
        L = C.CDLL(path, mode=os.RTLD_LOCAL)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)          # AttributeError if the library lacks a declared symbol
            f.restype = res
            f.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise AdvError(rc, lib().adv_last_error().decode(errors="replace"))
