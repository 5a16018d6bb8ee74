"""mcaat_b200 — B200-native (CUDA sm_100a) implementation of MCAAT's bounded-length multicycle search.

    gpu.py            ctypes binding of libmcaat_gpu.so (the C ABI in include/mcaat_gpu.h)
    cycle_finder.py   host-side mirror of the reference's Settings / CycleFinder interface
    csrc/             CUDA kernels + the C ABI
    host/             C++ drop-in `CycleFinder` for the reference's own host program
"""
from .gpu import DeviceGraph, MgError, lib, launch_count  # noqa: F401
from .cycle_finder import CycleFinder, CycleFinderSettings, Settings  # noqa: F401
