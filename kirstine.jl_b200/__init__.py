"""kirstine_b200: B200-native engine for Kirstine.jl's Bayesian design-objective hot path.

Layout: csrc/ (CUDA kernels + C ABI, built into libkirstine_b200.so), _ffi.py (ctypes binding of
include/kirstine_b200.h), api.py + designmeasure.py (host-side mirror of the reference's
user-facing API for this path).  No CPU fallback exists anywhere in this package.
"""
from . import _ffi
from ._ffi import EngineError, ProblemHandle, device_count, fp64_peak_probe
from .api import *  # noqa: F401,F403
from .api import _as_doptimality_problem  # noqa: F401
from .designmeasure import *  # noqa: F401,F403
from .sharding import shard_range  # noqa: F401
