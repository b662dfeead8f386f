"""The pure-C caller in examples/: the boundary is a C ABI, so a C99 program with nothing but the header and the
shared library must be able to drive the whole path (problem creation, swarm evaluation, information matrix, a
device-resident PSO run).  CPU: the example compiles as strict C99 against include/kirstine_b200.h, links, and
refuses to compute without a GPU.  GPU: its outputs match the oracle and the Python binding."""
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle as ko
from tests import workloads as wl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "kirstine.jl_b200")


def build_example(tmp_path):
    exe = str(tmp_path / "kst_objective_c")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "examples", "kst_objective_c.c"), "-L" + LIBDIR, "-lkirstine_b200",
           "-Wl,-rpath," + LIBDIR, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def write_input(path, model, crit, theta, pw, pts, w, mp, lb, ub, iters):
    S, K = w.shape
    with open(path, "wb") as f:
        f.write(struct.pack("<6i", model, crit, theta.shape[1], K, S, iters))
        f.write(struct.pack("<q", len(theta)))
        f.write(struct.pack("<3d", mp, lb, ub))
        for a in (theta, pw, pts, w):
            f.write(np.ascontiguousarray(a, dtype="<f8").tobytes())


def test_c_example_builds_and_refuses_without_gpu(tmp_path):
    import kirstine_b200 as kb
    exe = build_example(tmp_path)
    th = wl.sigemax_prior(10)
    pts, w = wl.random_swarm(4, 3, 0.0, 1.0, seed=1)
    write_input(tmp_path / "in.bin", 2, 0, th, np.full(10, 0.1), pts, w, 1.0, 0.0, 1.0, 0)
    if kb.device_count() == 0:
        r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
        assert r.returncode == 3 and "no CPU fallback" in r.stderr and not os.path.exists(tmp_path / "out.bin")


@pytest.mark.gpu
@pytest.mark.parametrize("model,crit,ub,K,prior", [(2, 0, 1.0, 8, "sig"), (3, 1, 48.0, 5, "tpc")])
def test_c_example_matches_oracle(tmp_path, model, crit, ub, K, prior):
    import kirstine_b200 as kb
    exe = build_example(tmp_path)
    th = wl.sigemax_prior(500, seed=3) if prior == "sig" else wl.tpc_prior(500, seed=3)
    pw = np.full(len(th), 1.0 / len(th))
    S, iters = 20, 5
    pts, w = wl.jittered_swarm(S, K, 0.0, ub, seed=2, log_lo=0.05 if prior == "tpc" else None)
    write_input(tmp_path / "in.bin", model, crit, th, pw, pts, w, 1.0, 0.0, ub, iters)
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = np.fromfile(tmp_path / "out.bin", dtype="<f8")
    rr = th.shape[1]
    fx, M, trace = out[:S], out[S:S + rr * rr].reshape(rr, rr), out[S + rr * rr:]
    o = ko.Problem(model, [1.0], th, pw, crit=crit)
    np.testing.assert_allclose(fx, o.objective_batch(pts, w), rtol=1e-10)
    np.testing.assert_allclose(M, o.informationmatrix(pts[0], w[0], 0), rtol=1e-10)
    # the same PSO through the Python binding: same seed, same prototypes -> the same trace, bit for bit
    h = kb.ProblemHandle(model, [1.0], th, pw, crit, None, (0.0,), (ub,))
    assert h.pso_init(S, K, pts, w, seed=12345) == 0
    want = [h.pso_state()["fg"]] + list(h.pso_run(iters - 1)[0])
    assert len(trace) == iters and np.array_equal(trace, np.array(want))
    assert np.all(np.diff(trace) >= 0) and trace[0] == fx.max()
