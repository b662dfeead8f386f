"""Dose-time-response family: test/example-dtr.jl, docs/src/dtr.md.
Five covariate parameterizations, r = 5, unit length up to 13, design regions of dimension 1 and 2."""
import numpy as np
import pytest

import oracle as ko
from tests import referee as rf
from tests import workloads as wl

pytestmark = pytest.mark.gpu
RTOL = 1e-10
kb = None

CASES = {  # name: (cp_id, cp_params, m, lb, ub)
    "copyboth": (2, [], 1, [0.0, 0.0], [100.0, 24.0]),
    "fixeddose": (3, [100.0], 1, [0.0], [24.0]),
    "fixedtimes": (4, list(np.arange(0.0, 26.0, 2.0)), 13, [0.0], [100.0]),
    "equidistant": (5, [], 13, [0.0, 0.0], [100.0, 2.0]),
    "logequidistant": (6, [np.sqrt(2.0)], 13, [0.0, 0.0], [100.0, 24 / 45.25483399593908]),
}


@pytest.fixture(scope="module", autouse=True)
def _engine():
    global kb
    import kirstine_b200
    kb = kirstine_b200
    assert kb.device_count() >= 1
    yield


def _pair(case, theta, crit=0, dt=None):
    cp, cpp, m, lb, ub = CASES[case]
    pw = np.full(len(theta), 1.0 / len(theta))
    h = kb.ProblemHandle(5, [1.0], theta, pw, crit, dt, lb, ub, m=m, cp_id=cp, cp_params=cpp)
    o = ko.Problem(ko.DTR, [1.0], theta, crit=crit, dt=dt, m=m, d=len(lb), cp_id=cp, cp_params=cpp)
    return h, o, (cp, cpp), m, lb, ub


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("crit", [0, 1])
def test_dtr_objective_vs_oracle(case, crit):
    theta = wl.dtr_prior(300)
    h, o, cov, m, lb, ub = _pair(case, theta, crit)
    K = 7 if m == 1 else 4
    pts, w = wl.jittered_swarm_nd(12, K, lb, ub, seed=3)
    pts[1, 0, 0] = 0.0                      # dose (or time) 0: trivial rows
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w, nthreads=8)
    pw = np.full(len(theta), 1.0 / len(theta))
    rf.assert_parity(got, want, lambda i: rf.objective(5, [1.0], pts[i], w[i], theta, pw, crit=crit, m=m, cov=cov),
                     rtol=RTOL, cond_fn=lambda i: rf.cond_max(5, [1.0], pts[i], w[i], theta, m, cov))
    assert np.isfinite(got).mean() >= 0.9
    M = h.informationmatrix(pts[0], w[0], 3)
    np.testing.assert_allclose(M, o.informationmatrix(pts[0], w[0], 3), rtol=1e-9, atol=1e-14 * np.abs(M).max())


def test_dtr_delta_method_and_gateaux():
    theta = wl.dtr_prior(200, seed=1)
    rng = np.random.default_rng(0)
    q = np.linalg.qr(rng.normal(size=(len(theta), 5, 5)))[0]
    dt = np.ascontiguousarray(np.transpose(q, (0, 2, 1))[:, :2, :])
    pw = np.full(len(theta), 1.0 / len(theta))
    for crit in (0, 1):
        h, o, cov, m, lb, ub = _pair("equidistant", theta, crit, dt)
        pts, w = wl.jittered_swarm_nd(6, 3, lb, ub, seed=5)
        got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
        rf.assert_parity(got, want, lambda i: rf.objective(5, [1.0], pts[i], w[i], theta, pw, crit=crit, dt=dt, m=m,
                                                           cov=cov), rtol=RTOL,
                         cond_fn=lambda i: rf.cond_max(5, [1.0], pts[i], w[i], theta, m, cov, dt=dt))
    for case in ("copyboth", "equidistant"):
        h, o, cov, m, lb, ub = _pair(case, theta)
        at_p, at_w = wl.jittered_swarm_nd(1, 6 if m == 1 else 3, lb, ub, seed=9)
        dirs, _ = wl.jittered_swarm_nd(1, 40, lb, ub, seed=10)
        assert not h.gateaux_prepare(at_p[0], at_w[0])
        got = h.gateaux_eval(dirs[0])
        want = o.gateauxderivative(at_p[0], at_w[0], dirs[0], nthreads=4)
        rf.assert_parity(got, want, lambda i: rf.gateaux(5, [1.0], at_p[0], at_w[0], [dirs[0][i]], theta, pw, m=m,
                                                         cov=cov)[0], rtol=RTOL, scale=np.abs(want).max())


def test_dtr_solve_and_equivalence_theorem():  # docs/src/dtr.md:263-300 (CopyBoth, Pso 30 x 100)
    prior = kb.PriorSample([kb.DTRParameter(*t) for t in wl.dtr_prior(500)])
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 100), time=(0, 24)),
                          model=kb.DTRMod(1, 1), covariate_parameterization=kb.CopyBoth(), prior_knowledge=prior)
    kb.seed(1357)
    proto = kb.DesignMeasure(*[([d, t], 1 / 6) for d, t in ((10, 2), (30, 4), (50, 8), (70, 12), (90, 16), (100, 24))])
    s, r = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(swarmsize=60, iterations=150), prototype=proto),
                    minposdose=1e-3, minpostime=1e-3, maxweight=1e-3, maxdist=2e-2)
    res = kb.optimization_result(r)
    assert np.all(np.diff(res.trace_fx) >= 0) and np.isfinite(res.maximum)
    grid = [kb.one_point_design([d, t]) for d in np.linspace(0, 100, 21) for t in np.linspace(0, 24, 25)]
    gd = kb.gateauxderivative(s, grid, dp)
    assert gd.max() < 0.15                      # D-efficiency >= exp(-0.15 / 5) = 0.97
    with pytest.raises(kb.EngineError, match="needs 2"):
        kb.objective(s, kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 100)),
                                         model=kb.DTRMod(1, 1), covariate_parameterization=kb.CopyBoth(),
                                         prior_knowledge=prior))


def test_dtr_maximum_unit_size():
    """K * m = KST_MAX_ROWS = 128 rows per particle with the widest row record: the particle tile plus the math
    tables need more than the 48 KB of shared memory a kernel gets by default (opt-in path)."""
    theta = wl.dtr_prior(100, seed=7)
    times = list(np.linspace(0.5, 24.0, 16))
    pw = np.full(len(theta), 1.0 / len(theta))
    h = kb.ProblemHandle(5, [1.0], theta, pw, 0, None, [0.0], [100.0], m=16, cp_id=4, cp_params=times)
    o = ko.Problem(ko.DTR, [1.0], theta, m=16, d=1, cp_id=4, cp_params=times)
    pts, w = wl.jittered_swarm_nd(9, 8, [0.0], [100.0], seed=11)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w, nthreads=8)
    rf.assert_parity(got, want, lambda i: rf.objective(5, [1.0], pts[i], w[i], theta, pw, m=16, cov=(4, times)),
                     rtol=RTOL, cond_fn=lambda i: rf.cond_max(5, [1.0], pts[i], w[i], theta, 16, (4, times)))
    M = h.informationmatrix(pts[0], w[0], 5)
    np.testing.assert_allclose(M, o.informationmatrix(pts[0], w[0], 5), rtol=1e-9, atol=1e-14 * np.abs(M).max())
    dirs = np.linspace(0.0, 100.0, 300)
    assert not h.gateaux_prepare(pts[0], w[0])
    gd, wgd = h.gateaux_eval(dirs), o.gateauxderivative(pts[0], w[0], dirs, nthreads=8)
    np.testing.assert_allclose(gd, wgd, rtol=1e-8, atol=1e-8 * np.abs(wgd).max())
    with pytest.raises(kb.EngineError):
        h.objective_batch(np.zeros((2, 9, 1)), np.full((2, 9), 1 / 9))   # 9 * 16 rows > KST_MAX_ROWS
