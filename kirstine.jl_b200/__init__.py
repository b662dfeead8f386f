"""kirstine_b200: B200-native engine for Kirstine.jl's Bayesian design-objective hot path.

Layout: csrc/ (CUDA kernels + C ABI, built into libkirstine_b200.so), _ffi.py (ctypes binding of
include/kirstine_b200.h), and the host-side mirror of the reference's user-facing API for this path:
models.py (registered families), designmeasure.py, problem.py (DesignProblem and the public evaluators),
solve.py (optimize, solve, the two strategies).

The mirror keeps the reference's names, argument meaning and error behaviour (exported list:
src/Kirstine.jl:25-96), so parity tests read like the reference's own tests.  Everything that
evaluates an objective or a Gateaux derivative goes through the C ABI (include/kirstine_b200.h)
to the CUDA engine; there is no CPU fallback anywhere in this package.  Differences forced by the host
language: indices are 0-based, keyword structs become Python classes, `DeltaMethod` takes a Python callable
that is evaluated on the host once per draw (exactly what trafo_constants does,
src/transformation.jl:71-92), and only *registered* model families (csrc/kst_models.cuh) can be
used -- arbitrary user `jacobianmatrix!` code cannot run on the GPU.
"""
from . import _ffi
from ._ffi import EngineError, ProblemHandle, device_count, fp64_peak_probe
from .designmeasure import *  # noqa: F401,F403
from .models import *  # noqa: F401,F403
from .problem import *  # noqa: F401,F403
from .problem import _as_doptimality_problem  # noqa: F401
from .sharding import shard_range  # noqa: F401
from .solve import *  # noqa: F401,F403
from .solve import _solve_with_direct, _solve_with_exchange  # noqa: F401
