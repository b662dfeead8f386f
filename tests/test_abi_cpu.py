"""CPU-side checks of the C-ABI library: it loads without a GPU, exports every symbol that
include/kirstine_b200.h declares, and refuses to compute without a CUDA device (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kirstine_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"KST_API[^;(]*?\b(kst_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import ctypes

    import kirstine_b200 as kb
    names = _declared_symbols()
    assert len(names) >= 20
    lib = ctypes.CDLL(kb._ffi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # the ctypes binding covers exactly the declared surface
    assert sorted(kb._ffi.SIGNATURES) == names
    assert kb._ffi.lib().kst_version() == 100


def test_no_cpu_fallback_without_device():
    import kirstine_b200 as kb
    if kb.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(kb.EngineError, match="no CUDA device"):
        kb.ProblemHandle(1, [1.0], [[1.0, 10.0, 5.0]], [1.0], 0, None, [0.0], [10.0])
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 10)), model=kb.EmaxModel(1),
                          covariate_parameterization=kb.CopyDose(),
                          prior_knowledge=kb.PriorSample([kb.EmaxPar(1, 10, 5)]))
    with pytest.raises(kb.EngineError):
        kb.objective(kb.uniform_design([[0], [2.5], [10]]), dp)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kirstine.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower() or f == "kst_pso.cu" and "import oracle" not in src, f
