"""opencl_path_tracer_b200 — the prog.cl path-trace hot path of zotya701/OpenCL_Path_tracer on B200.

The product is ``libptb200.so`` (CUDA, sm_100a, C-ABI in ``include/ptb200.h``).  This package is the
thin Python host around it: ctypes bindings (:mod:`capi`), a mirror of the reference host's ``Scene``
interface (:mod:`host`), procedural scenes (:mod:`scenes`) and the multi-GPU drivers (:mod:`multigpu`).
There is no CPU fallback anywhere in this package: if the CUDA library is missing or no GPU is
present, the calls raise.
"""
from .capi import Ptb, PtbError, lib_path, load_library  # noqa: F401

__all__ = ["Ptb", "PtbError", "lib_path", "load_library"]
