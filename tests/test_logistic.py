"""Logistic regression, the user-defined Model supertype of docs/src/extend-model.md:59-106, with the two concrete
families of that vignette (LinLogReg :123-179, SinLogReg :330-355).

Known answers: the three solved designs the vignette asserts (s0 :216-219 = Fedorov & Leonov 2013 p. 154,
s1 :285-289 with the odds-ratio DeltaMethod, s2 :388-392).  CPU tests pin the oracle to them and to a plain numpy
restatement; GPU tests check the engine against the oracle and re-solve the problems.
"""
import numpy as np
import pytest

import oracle as ko

RTOL = 1e-10

S0 = ([[0.8047475621673699], [2.5101732141323296]], [0.49999389060536853, 0.5000061093946315])
S1 = ([[0.34921961868772716, 0.0], [0.9841082624551549, 0.0], [1.0, 1.0]],
      [0.3543739967246416, 0.35433329676914754, 0.29129270650621086])
S2 = ([[0.10724768400759671], [0.8412520842064646], [1.0]],
      [0.3335078264280175, 0.33343038623171556, 0.33306178734026687])
TH0, TH1, TH2 = [[-3.0, 1.81]], [[-4.0, 6.0, -3.0]], [[1.0, 1.5, 0.125]]


def odds_ratio_dt(theta):  # docs/src/extend-model.md:274: p -> [0 exp(b1) 0; 0 0 exp(b2)]
    return np.stack([np.array([[0, np.exp(t[1]), 0], [0, 0, np.exp(t[2])]]) for t in np.asarray(theta, float)])


def grid2(n):
    g = np.linspace(0, 1, n)
    return np.stack(np.meshgrid(g, g), -1).reshape(-1, 2)


def llr_prior(n, d, seed):
    rng = np.random.default_rng(seed)
    return np.column_stack([rng.normal(-1.0, 0.5, n)] + [rng.normal(1.5, 0.7, n) for _ in range(d)])


def slr_prior(n, seed):
    rng = np.random.default_rng(seed)
    return np.column_stack([rng.uniform(0.5, 1.5, n), rng.uniform(1.0, 2.0, n), rng.uniform(0.0, 0.5, n)])


# ---- CPU: the oracle -------------------------------------------------------------------------------------------

def test_oracle_logistic_informationmatrix_vs_numpy():
    th = llr_prior(3, 2, 1)
    o = ko.Problem(ko.LINLOGREG, [], th, d=2)
    pts, w = np.array(S1[0]), np.array(S1[1])
    for j in range(3):
        M = np.zeros((3, 3))
        for x, wk in zip(pts, w):
            jm = np.array([1.0, x[0], x[1]])
            p = 1 / (1 + np.exp(-(th[j] @ jm)))
            M += p * (1 - p) * wk * np.outer(jm, jm)
        np.testing.assert_allclose(o.informationmatrix(pts, w, j), M, rtol=1e-14)
    th = slr_prior(2, 2)
    o = ko.Problem(ko.SINLOGREG, [], th, d=1)
    a, b, c = th[1]
    M = np.zeros((3, 3))
    for (t,), wk in zip(S2[0], S2[1]):
        ph = 2 * np.pi * b * (t - c)
        jm = np.array([np.sin(ph), a * np.cos(ph) * 2 * np.pi * (t - c), -a * np.cos(ph) * 2 * np.pi * b])
        p = 1 / (1 + np.exp(-a * np.sin(ph)))
        M += p * (1 - p) * wk * np.outer(jm, jm)
    np.testing.assert_allclose(o.informationmatrix(S2[0], S2[1], 1), M, rtol=1e-13)


def test_oracle_logistic_golden_designs_satisfy_equivalence_theorem():
    o0 = ko.Problem(ko.LINLOGREG, [], TH0, d=1)
    g0 = o0.gateauxderivative(S0[0], S0[1], np.linspace(0, 4, 401))
    assert g0.max() < 1e-4 and np.all(np.abs(o0.gateauxderivative(S0[0], S0[1], S0[0])) < 1e-4)
    o1 = ko.Problem(ko.LINLOGREG, [], TH1, d=2, dt=odds_ratio_dt(TH1))
    assert o1.gateauxderivative(S1[0], S1[1], grid2(51)).max() < 1e-3
    o2 = ko.Problem(ko.SINLOGREG, [], TH2, d=1)
    assert o2.gateauxderivative(S2[0], S2[1], np.linspace(0, 1, 401)).max() < 5e-3


# ---- GPU: the engine ---------------------------------------------------------------------------------------------

def H(model, theta, d, lb, ub, crit=0, dt=None, pw=None):
    import kirstine_b200 as kb
    pw = np.full(len(theta), 1.0 / len(theta)) if pw is None else pw
    return kb.ProblemHandle(model, [0.0], theta, pw, crit, dt, lb, ub)


def _swarm(S, K, d, lo, hi, seed):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(lo, hi, (S, K, d))
    w = rng.dirichlet(np.full(K, 4.0), S)
    return pts, w


@pytest.mark.gpu
@pytest.mark.parametrize("d", [1, 2, 3])
@pytest.mark.parametrize("crit", [0, 1, 2])
def test_engine_linlogreg_vs_oracle(d, crit):
    th = llr_prior(200, d, 10 + d)
    pw = np.random.default_rng(4).dirichlet(np.ones(len(th)))
    h, o = H(6, th, d, (0.0,) * d, (1.0,) * d, crit, pw=pw), ko.Problem(ko.LINLOGREG, [], th, pw, crit=crit, d=d)
    pts, w = _swarm(30, d + 3, d, 0.0, 1.0, 3)
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)
    np.testing.assert_allclose(h.informationmatrix(pts[2], w[2], 17), o.informationmatrix(pts[2], w[2], 17),
                               rtol=RTOL, atol=1e-16)
    dirs = np.random.default_rng(5).uniform(0, 1, (300, d))
    assert not h.gateaux_prepare(pts[1], w[1])
    want = o.gateauxderivative(pts[1], w[1], dirs)
    np.testing.assert_allclose(h.gateaux_eval(dirs), want, rtol=RTOL, atol=RTOL * np.abs(want).max())


@pytest.mark.gpu
def test_engine_linlogreg_delta_and_singular():
    th = llr_prior(150, 2, 21)
    dt = odds_ratio_dt(th)
    for crit in (0, 1):
        h, o = H(6, th, 2, (0.0, 0.0), (1.0, 1.0), crit, dt), ko.Problem(ko.LINLOGREG, [], th, crit=crit, d=2, dt=dt)
        pts, w = _swarm(20, 5, 2, 0.0, 1.0, 8)
        np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)
    h = H(6, th, 2, (0.0, 0.0), (1.0, 1.0))
    assert h.objective_batch([[[0.2, 0.3], [0.7, 0.1]]], [[0.5, 0.5]])[0] == -np.inf   # 2 points, 3 parameters
    # the covariate dimension has to match the region, and only d = 1..3 is registered
    import kirstine_b200 as kb
    with pytest.raises(kb.EngineError):
        kb.ProblemHandle(6, [0.0], llr_prior(5, 2, 1), np.full(5, 0.2), 0, None, (0.0,), (1.0,))
    with pytest.raises(kb.EngineError):
        kb.ProblemHandle(6, [0.0], llr_prior(5, 4, 1), np.full(5, 0.2), 0, None, (0.0,) * 4, (1.0,) * 4)


@pytest.mark.gpu
@pytest.mark.parametrize("crit", [0, 1])
def test_engine_sinlogreg_vs_oracle(crit):
    th = slr_prior(300, 6)
    h, o = H(7, th, 1, (0.0,), (1.0,), crit), ko.Problem(ko.SINLOGREG, [], th, crit=crit, d=1)
    pts, w = _swarm(30, 5, 1, 0.0, 1.0, 9)
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)
    dirs = np.linspace(0, 1, 201)
    assert not h.gateaux_prepare(pts[0], w[0])
    want = o.gateauxderivative(pts[0], w[0], dirs)
    np.testing.assert_allclose(h.gateaux_eval(dirs), want, rtol=RTOL, atol=RTOL * np.abs(want).max())


@pytest.mark.gpu
def test_engine_logistic_extreme_predictor_rows_vanish():
    # |eta| far beyond the exponent range: p (1 - p) underflows to 0 in the reference, the row contributes nothing
    th = np.array([[-3.0, 1.81]])
    h, o = H(6, th, 1, (0.0,), (4000.0,)), ko.Problem(ko.LINLOGREG, [], th, d=1)
    pts, w = [[[0.8], [2.5], [3000.0]]], [[0.4, 0.4, 0.2]]
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)


@pytest.mark.gpu
def test_engine_solves_logistic_vignette_problems():
    """docs/src/extend-model.md:196-219, 266-289, 368-392 with a larger swarm than the vignette (other RNG stream)."""
    import kirstine_b200 as kb
    kb.seed(12345)
    pso = kb.Pso(swarmsize=200, iterations=200)
    dp0 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(("x1", (0, 4))),
                           covariate_parameterization=kb.CopyVector(),
                           prior_knowledge=kb.PriorSample([kb.LLRParameter(-3, [1.81])]), model=kb.LinLogReg(1))
    s0, _ = kb.solve(dp0, kb.DirectMaximization(optimizer=pso, prototype=kb.random_design(kb.region(dp0), 4)),
                     maxdist=1e-3)
    assert kb.efficiency(s0, kb.DesignMeasure(S0[0], S0[1]), dp0) >= 0.9999
    dp1 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(("x1", (0, 1)), ("x2", (0, 1))),
                           covariate_parameterization=kb.CopyVector(),
                           prior_knowledge=kb.PriorSample([kb.LLRParameter(-4, [6, -3])]), model=kb.LinLogReg(2),
                           transformation=kb.DeltaMethod(lambda p: np.array(
                               [[0, np.exp(p.beta_rest[0]), 0], [0, 0, np.exp(p.beta_rest[1])]])))
    s1, _ = kb.solve(dp1, kb.DirectMaximization(optimizer=pso, prototype=kb.random_design(kb.region(dp1), 4)),
                     maxdist=5e-2, maxweight=1e-3)
    assert kb.efficiency(s1, kb.DesignMeasure(S1[0], S1[1]), dp1) >= 0.9999
    dp2 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(("t", (0, 1))),
                           covariate_parameterization=kb.JustCopy("t"),
                           prior_knowledge=kb.PriorSample([kb.SLRParameter(a=1, b=1.5, c=0.125)]),
                           model=kb.SinLogReg())
    s2, _ = kb.solve(dp2, kb.DirectMaximization(optimizer=pso, prototype=kb.random_design(kb.region(dp2), 4)),
                     maxdist=5e-2, maxweight=1e-3)
    assert kb.efficiency(s2, kb.DesignMeasure(S2[0], S2[1]), dp2) >= 0.9999
