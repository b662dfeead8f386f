"""kst_exchange_run (the Exchange strategy of src/solve-exchange.jl:116-179 behind one C-ABI call) against the same loop
stepped from the host, against the oracle on the shared random stream, and against the equivalence theorem."""
import numpy as np
import pytest

import oracle as ko
from tests import workloads as wl

pytestmark = pytest.mark.gpu
kb = None


@pytest.fixture(scope="module", autouse=True)
def _engine():
    global kb
    import kirstine_b200
    kb = kirstine_b200
    assert kb.device_count() >= 1
    yield


def _tpc_problem(n=50, seed=1):
    prior = kb.PriorSample([kb.TPCParameter(*t) for t in wl.tpc_prior(n, seed=seed)])
    return kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(time=(0, 48)), model=kb.TPCModel(sigma=1),
                            covariate_parameterization=kb.CopyTime(), prior_knowledge=prior)


def test_exchange_device_call_equals_host_stepped_loop():
    """trace_state = False runs kst_exchange_run; trace_state = True steps the same loop from Python, one
    kst_pso_init + kst_pso_run(1) per iteration.  Same seeds, same kernels: identical designs and traces."""
    dp = _tpc_problem()
    cand = kb.DesignMeasure(([0.5], 0.4), ([3.0], 0.3), ([30.0], 0.3))
    st = kb.Exchange(optimizer_direction=kb.Pso(iterations=8, swarmsize=25), optimizer_weight=kb.Pso(iterations=9, swarmsize=20),
                     steps=5, candidate=cand, simplify_args=dict(maxweight=1e-4, maxdist=2e-3))
    kb.seed(99)
    d1, r1 = kb.solve(dp, st)
    kb.seed(99)
    d2, r2 = kb.solve(dp, st, trace_state=True)
    assert np.array_equal(d1.points, d2.points) and np.array_equal(d1.weights, d2.weights)
    assert len(r1.ord) == len(r1.orw) == 5
    for a, b in zip(r1.ord + r1.orw, r2.ord + r2.orw):
        assert a.trace_fx == b.trace_fx and a.maximum == b.maximum and a.n_eval == b.n_eval
        assert all(x == y for x, y in zip(a.trace_x, b.trace_x))
        assert a.maximizer == b.maximizer
        assert len(a.trace_state) == 1 and len(b.trace_state) == len(b.trace_fx)
        s0a, s0b = a.trace_state[0], b.trace_state[0]
        for key in ("x_pts", "x_w", "p_pts", "p_w", "v_pts", "v_w", "g_pts", "g_w", "fx", "fp"):
            assert np.array_equal(s0a.raw[key], s0b.raw[key], equal_nan=True), key
        assert s0a.fg == s0b.fg
    assert all(kb.numpoints(o.maximizer) == 1 for o in r1.ord)
    assert all(np.all(np.diff(o.trace_fx) >= 0) for o in r1.ord + r1.orw)


def test_one_exchange_step_against_oracle():
    """One step -- direction search, prepended point, reweighting -- replayed with the oracle's swarm on the shared
    Philox stream (seeds: the call's seed for the direction run, its LCG successor for the weight run)."""
    th = wl.tpc_prior(120, seed=6)
    pw = np.full(len(th), 1.0 / len(th))
    lb, ub = [0.0], [48.0]
    h = kb.ProblemHandle(3, [1.0], th, pw, 0, None, lb, ub)
    o = ko.Problem(ko.TPC, [1.0], th, pw)
    cand_p, cand_w = np.array([[0.4], [2.5], [25.0]]), np.array([0.3, 0.3, 0.4])
    Sd, itd, Sw, itw, seed = 20, 6, 16, 7, 4242
    r = h.exchange_run(cand_p, cand_w, 1, (Sd, itd, 2.05, 2.05), (Sw, itw, 2.05, 2.05), seed=seed)
    # direction run
    gconst = o.gateaux_constants(cand_p, cand_w)
    po = ko.Pso(o, Sd, 1, lb, ub, seed=seed, mode=1, gconst=gconst)
    po.init(cand_p.reshape(3, 1, 1), np.ones((3, 1)))
    want_fx = [po.state()["fg"]]
    for _ in range(itd - 1):
        po.tick()
        want_fx.append(po.state()["fg"])
    direction = po.state()["g_pts"][0]
    np.testing.assert_allclose(r["dir_fx"][0], want_fx, rtol=1e-9, atol=1e-10 * np.abs(want_fx).max())
    np.testing.assert_allclose(r["dir_x"][0, -1], direction, rtol=1e-11)
    # the new point goes to the front with weight 1 / (1 + K); the others shrink by K / (1 + K)
    K = 3
    mix_p = np.vstack([r["dir_x"][0, -1].reshape(1, 1), cand_p])
    mix_w = np.concatenate([[1 / (1 + K)], (1 - 1 / (1 + K)) * cand_w])
    s0 = r["w_state0"][0]
    assert np.array_equal(s0["x_pts"][0], mix_p) and np.array_equal(s0["x_w"][0], mix_w)
    # weight run: all points fixed
    seed2 = (seed * 6364136223846793005 + 1442695040888963407) % (1 << 64)
    pw_ = ko.Pso(o, Sw, K + 1, lb, ub, fixw=[0] * (K + 1), fixp=[1] * (K + 1), seed=seed2)
    pw_.init(mix_p.reshape(1, K + 1, 1), mix_w.reshape(1, K + 1))
    want = [pw_.state()["fg"]]
    for _ in range(itw - 1):
        pw_.tick()
        want.append(pw_.state()["fg"])
    np.testing.assert_allclose(r["w_fx"][0], want, rtol=1e-10)
    fin = pw_.state()
    np.testing.assert_allclose(r["weights"], fin["g_w"], rtol=1e-9, atol=1e-12)
    assert np.array_equal(r["points"], mix_p) and r["w_K"][0] == K + 1


def test_exchange_reaches_the_optimal_design():
    """The north-star bar for solved designs: >= 0.9999 efficiency against direct maximisation, and the equivalence
    theorem at the solution (test/solve-exchange.jl:69-88 checks structure only; the bar is BASELINE.json's)."""
    dp = _tpc_problem()
    kb.seed(1357)
    cand = kb.DesignMeasure(([0.5], 0.4), ([3.0], 0.3), ([30.0], 0.3))
    str_ = kb.Exchange(optimizer_direction=kb.Pso(iterations=40, swarmsize=50),
                       optimizer_weight=kb.Pso(iterations=60, swarmsize=50), steps=15, candidate=cand,
                       simplify_args=dict(maxweight=1e-4, maxdist=2e-3))
    d, r = kb.solve(dp, str_, maxweight=1e-3, maxdist=1e-2)
    assert len(r.ord) == len(r.orw) == 15
    assert kb.objective(d, dp) > kb.objective(cand, dp)
    grid = np.exp(np.linspace(np.log(0.05), np.log(48), 400))
    gd = kb.gateauxderivative(d, [kb.one_point_design([t]) for t in grid], dp)
    assert gd.max() <= 1e-3, gd.max()
    kb.seed(2468)
    dm, _ = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(iterations=200, swarmsize=100),
                                               prototype=kb.DesignMeasure(([0.3], 1 / 3), ([2.0], 1 / 3), ([20.0], 1 / 3))))
    assert kb.efficiency(d, dm, dp) >= 0.9999


def test_exchange_errors_and_dtr_simplify_unique():
    th = wl.tpc_prior(20, seed=2)
    h = kb.ProblemHandle(3, [1.0], th, np.full(20, 0.05), 0, None, [0.0], [48.0])
    # a one-point candidate is singular for three parameters: the reference throws SingularException
    with pytest.raises(kb.EngineError) as ei:
        h.exchange_run([[1.0]], [1.0], 2, (10, 3, 2.05, 2.05), (10, 3, 2.05, 2.05))
    assert ei.value.code == -7
    with pytest.raises(kb.EngineError, match="swarmsize must be >= number of prototype"):
        h.exchange_run([[1.0], [2.0], [9.0]], [0.3, 0.3, 0.4], 1, (2, 3, 2.05, 2.05), (10, 3, 2.05, 2.05))
    with pytest.raises(kb.EngineError, match="outside design region"):
        h.exchange_run([[1.0], [2.0], [99.0]], [0.3, 0.3, 0.4], 1, (10, 3, 2.05, 2.05), (10, 3, 2.05, 2.05))
    # dose-time-response with CopyBoth: the covariate parameterization's own simplify_unique (minposdose / minpostime)
    # runs inside the call and must give what the host-stepped loop gives
    prior = kb.PriorSample([kb.DTRParameter(*t) for t in wl.dtr_prior(100)])
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 100), time=(0, 24)),
                          model=kb.DTRMod(1, 1), covariate_parameterization=kb.CopyBoth(), prior_knowledge=prior)
    cand = kb.DesignMeasure(*[([d_, t], 1 / 6) for d_, t in ((10, 2), (30, 4), (50, 8), (70, 12), (90, 16), (100, 24))])
    st = kb.Exchange(optimizer_direction=kb.Pso(iterations=6, swarmsize=20), optimizer_weight=kb.Pso(iterations=6, swarmsize=20),
                     steps=3, candidate=cand, simplify_args=dict(maxweight=1e-3, maxdist=1e-2, minposdose=1e-3, minpostime=1e-3))
    kb.seed(5)
    d1, _ = kb.solve(dp, st)
    kb.seed(5)
    d2, _ = kb.solve(dp, st, trace_state=True)
    assert d1 == d2
