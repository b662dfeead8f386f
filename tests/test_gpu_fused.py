"""The persistent swarm kernel (csrc/kst_pso_fused.cuh) against the three-launch path it replaces for small
problems, and kst_pso_optimize against kst_pso_init + kst_pso_run.  Both paths run the same device functions on the
same grid, so everything must agree bit for bit; the oracle comparisons of the swarm live in test_gpu_pso.py and
run through whichever path the problem size selects (KST_NO_FUSED=1 forces the three-launch path)."""
import os

import numpy as np
import pytest

from tests import workloads as wl

pytestmark = pytest.mark.gpu
kb = None

KEYS = ("x_pts", "x_w", "p_pts", "p_w", "v_pts", "v_w", "g_pts", "g_w", "fx", "fp", "fg", "n_eval")


@pytest.fixture(scope="module", autouse=True)
def _engine():
    global kb
    import kirstine_b200
    kb = kirstine_b200
    assert kb.device_count() >= 1
    yield
    os.environ.pop("KST_NO_FUSED", None)


def _unfused(on):
    if on:
        os.environ["KST_NO_FUSED"] = "1"
    else:
        os.environ.pop("KST_NO_FUSED", None)


def _same(a, b, what):
    for k in KEYS:
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k]), equal_nan=True), "%s: %s differs" % (what, k)


CASES = {
    # name: (model, mp, prior, ub, K, S, crit, delta, m, cp, fixw, fixp)
    "cfg1 emax": (1, [1.0], lambda: wl.emax_prior(1000), 10.0, 3, 50, 0, False, 1, 0, None, None),
    "cfg1b sigemax": (2, [1.0], lambda: wl.sigemax_prior(1000), 1.0, 8, 100, 0, False, 1, 0, None, None),
    "cfg2 A delta": (2, [1.0], lambda: wl.sigemax_prior(4000, seed=2), 1.0, 6, 64, 1, True, 1, 0, None, None),
    "tpc nonuniform S=7": (3, [1.0], lambda: wl.tpc_prior(300), 48.0, 4, 7, 0, False, 1, 0, None, None),
    "vu fixed": (4, [0.1], lambda: wl.tpc_prior(500, seed=4), 24.0, 3, 25, 0, False, 3, 1, [0, 0, 0], [1, 1, 1]),
    "emax fixed w": (1, [1.0], lambda: wl.emax_prior(200), 10.0, 3, 33, 1, False, 1, 0, [1, 0, 0], [0, 0, 1]),
}


def _handle(case):
    model, mp, prior, ub, K, S, crit, delta, m, cp, fixw, fixp = CASES[case]
    th = prior()
    n = len(th)
    pw = np.full(n, 1.0 / n)
    if "nonuniform" in case:
        pw = np.random.default_rng(1).dirichlet(np.ones(n))
    dt = wl.ded90_jacobians(th) if delta else None
    h = kb.ProblemHandle(model, mp, th, pw, crit, dt, [0.0], [ub], m=m, cp_id=cp)
    lo = 0.05 * ub
    proto = (np.linspace(lo, ub, K).reshape(K, 1), np.full(K, 1.0 / K))
    return h, S, K, proto, fixw, fixp


@pytest.mark.parametrize("case", list(CASES))
def test_fused_run_equals_three_launch_path(case):
    res = {}
    for unf in (True, False):
        _unfused(unf)
        h, S, K, proto, fixw, fixp = _handle(case)
        assert h.pso_init(S, K, *proto, fixw=fixw, fixp=fixp, seed=77) == 0
        tr = [h.pso_run(5), h.pso_run(1), h.pso_run(6)]     # odd and even run lengths: both buffer parities
        res[unf] = (h.pso_state(), tr)
    _unfused(False)
    _same(res[True][0], res[False][0], case)
    for a, b in zip(res[True][1], res[False][1]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y), case
    assert np.all(np.diff(np.concatenate([t[0] for t in res[False][1]])) >= 0)


@pytest.mark.parametrize("case", list(CASES))
def test_pso_optimize_equals_init_plus_run(case):
    iters = 9
    h, S, K, proto, fixw, fixp = _handle(case)
    assert h.pso_init(S, K, *proto, fixw=fixw, fixp=fixp, seed=5) == 0
    s0 = h.pso_state()
    tfx, tp, tw = h.pso_run(iters - 1)
    final = h.pso_state()
    for unf in (False, True):
        _unfused(unf)
        h2, *_ = _handle(case)
        st, fx, px, pw, snap = h2.pso_optimize(S, K, iters, *proto, fixw=fixw, fixp=fixp, seed=5)
        assert st == 0
        assert fx[0] == s0["fg"] and np.array_equal(px[0], s0["g_pts"]) and np.array_equal(pw[0], s0["g_w"])
        assert np.array_equal(fx[1:], tfx) and np.array_equal(px[1:], tp) and np.array_equal(pw[1:], tw)
        _same(snap, s0, "%s initial state (unfused=%s)" % (case, unf))
        fin2 = h2.pso_state()
        _same(fin2, final, "%s final state (unfused=%s)" % (case, unf))
        assert fin2["n_eval"] == S * iters
        # the swarm carries on from where optimize left it
        a = h2.pso_run(2)
    _unfused(False)
    b = h.pso_run(2)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_gateaux_mode_fused_equals_three_launch_path_and_grid_kernel():
    th = wl.tpc_prior(1000, seed=4)
    pw = np.full(len(th), 1e-3)
    at_p, at_w = [2.0, 8.0, 20.0], [0.3, 0.3, 0.4]
    out = {}
    for unf in (True, False):
        _unfused(unf)
        h = kb.ProblemHandle(4, [0.1], th, pw, 0, None, [0.0], [24.0], m=3, cp_id=1)
        assert not h.gateaux_prepare(at_p, at_w)
        st, fx, px, pwt, snap = h.pso_optimize(25, 1, 10, np.reshape(at_p, (3, 1, 1)), np.ones((3, 1)), seed=3,
                                               mode=kb._ffi.PSO_MODE_GATEAUX)
        assert st == 0
        out[unf] = (fx, px, h.pso_state(), snap)
        # the swarm's values are the Gateaux derivative the grid kernel computes for the same directions
        state = out[unf][2]
        gd = h.gateaux_eval(state["x_pts"].reshape(-1, 1))
        np.testing.assert_allclose(state["fx"], gd, rtol=1e-11, atol=1e-11 * np.abs(gd).max())
    _unfused(False)
    assert np.array_equal(out[True][0], out[False][0]) and np.array_equal(out[True][1], out[False][1])
    _same(out[True][2], out[False][2], "gateaux final")
    _same(out[True][3], out[False][3], "gateaux initial")
    assert np.all(np.diff(out[False][0]) >= 0)
    # singular design: every value NaN, status 2 (NaN on all initial points), on both paths
    for unf in (True, False):
        _unfused(unf)
        h = kb.ProblemHandle(4, [0.1], th, pw, 0, None, [0.0], [24.0], m=1, cp_id=1)
        assert h.gateaux_prepare([1.0], [1.0])
        st, *_ = h.pso_optimize(8, 1, 3, [[[1.0]]], [[1.0]], mode=kb._ffi.PSO_MODE_GATEAUX)
        assert st == 2
    _unfused(False)


def test_status_words_through_optimize():
    theta = np.array([[1.0, 10, 5]])
    h = kb.ProblemHandle(1, [1.0], theta, [float("nan")], 0, None, [0.0], [10.0])
    st, *_ = h.pso_optimize(5, 3, 4, [[0.0], [5.0], [10.0]], [1 / 3] * 3)
    assert st == 2
    with pytest.raises(kb.EngineError, match="iterations must be positive"):
        h.pso_optimize(5, 3, 0, [[0.0], [5.0], [10.0]], [1 / 3] * 3)
    with pytest.raises(kb.EngineError, match="at least 2"):
        h.pso_optimize(1, 3, 4, [[0.0], [5.0], [10.0]], [1 / 3] * 3)


def test_large_swarm_falls_back_and_matches():
    """A problem beyond one wave takes the three-launch path inside kst_pso_optimize; same contract."""
    th = wl.sigemax_prior(60_000, seed=5)
    h = kb.ProblemHandle(2, [1.0], th, np.full(len(th), 1.0 / len(th)), 0, None, [0.0], [1.0])
    proto = (np.linspace(0, 1, 8).reshape(8, 1), np.full(8, 0.125))
    st, fx, px, pw, snap = h.pso_optimize(512, 8, 4, *proto, seed=9)
    assert st == 0 and np.all(np.diff(fx) >= 0)
    h2 = kb.ProblemHandle(2, [1.0], th, np.full(len(th), 1.0 / len(th)), 0, None, [0.0], [1.0])
    assert h2.pso_init(512, 8, *proto, seed=9) == 0
    _same(snap, h2.pso_state(), "large initial")
    t2 = h2.pso_run(3)
    assert np.array_equal(fx[1:], t2[0])
    _same(h.pso_state(), h2.pso_state(), "large final")
