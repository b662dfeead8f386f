"""Host-side mirror of Kirstine.jl's user-facing API for the accelerated path.

Same names, argument meaning and error behaviour as the reference (exported list:
src/Kirstine.jl:25-96), so parity tests read like the reference's own tests.  Everything that
evaluates an objective or a Gateaux derivative goes through the C ABI (include/kirstine_b200.h)
to the CUDA engine; there is no CPU fallback.  Differences forced by the host language:
indices are 0-based, keyword structs become Python classes, `DeltaMethod` takes a Python callable
that is evaluated on the host once per draw (exactly what trafo_constants does,
src/transformation.jl:71-92), and only *registered* model families (csrc/kst_models.cuh) can be
used -- arbitrary user `jacobianmatrix!` code cannot run on the GPU.
"""
from .models import *  # noqa: F401,F403
from .problem import *  # noqa: F401,F403
from .problem import _as_doptimality_problem  # noqa: F401
from .solve import *  # noqa: F401,F403
from .solve import _solve_with_direct, _solve_with_exchange  # noqa: F401
