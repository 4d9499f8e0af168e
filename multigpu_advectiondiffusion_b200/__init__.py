"""B200-native explicit finite-difference time stepping (heat 2-D/3-D, Burgers 3-D WENO5,
SSP-RK3) -- host-side Python mirror of the C ABI in include/advdiff.h.

The compute path is libadvdiff.so (hand-written sm_100a CUDA, csrc/).  There is no CPU or
PyTorch fallback: importing works anywhere (so the ABI can be inspected), but creating a
solver without the built library or without a B200 raises.
"""
from ._lib import (AdvError, Desc, StageArgs, TARGETS, F32, F64, FLAG_VISC_ONCE, lib, lib_path,
                   build_library)
from .solver import Solver, solve_host, reference_tEnd_for_iters

__all__ = ["AdvError", "Desc", "StageArgs", "TARGETS", "F32", "F64", "FLAG_VISC_ONCE", "lib",
           "lib_path", "build_library", "Solver", "solve_host", "reference_tEnd_for_iters"]
