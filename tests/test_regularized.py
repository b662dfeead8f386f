"""RegularizedFisherMatrix, the user-extension NormalApproximation of docs/src/extend-approximation.md:69-79:
informationmatrix!(afm, na) adds a constant matrix R to every average Fisher matrix (objective, Gateaux constants
and the direction's matrix alike).

Known answer: the vignette's locally optimal design for the time to maximum concentration (s_tr, :171-176), which
is singular without R and solvable with R = 1e-4 I.
"""
import numpy as np
import pytest

import oracle as ko
from tests import workloads as wl

RTOL = 1e-10
TPC_G1 = [4.298, 0.05884, 21.80]
S_TR = ([0.18009461306034247, 3.561491759806154], [0.6051847764968658, 0.39481522350313425])
R4 = np.eye(3) * 1e-4


def dttm(p):  # docs/src/extend-approximation.md:103-111
    a, e, _ = p
    A = a - e
    B = np.log(a / e)
    return np.array([[(A / a - B) / A ** 2, (B - A / e) / A ** 2, 0.0]])


def spd(r, seed, scale):
    q = np.random.default_rng(seed).normal(size=(r, r))
    return scale * (q @ q.T + r * np.eye(r))


# ---- CPU: the oracle -------------------------------------------------------------------------------------------

def test_oracle_regularized_informationmatrix_and_objective():
    th = wl.tpc_prior(5, seed=3)
    R = spd(3, 1, 1e-3)
    o, oreg = ko.Problem(ko.TPC, [1.0], th), ko.Problem(ko.TPC, [1.0], th).set_regularization(R)
    pts, w = [0.5, 2.0, 10.0, 30.0], [0.1, 0.2, 0.3, 0.4]
    for j in range(5):
        np.testing.assert_allclose(oreg.informationmatrix(pts, w, j), o.informationmatrix(pts, w, j) + R, rtol=1e-15)
    want = np.mean([np.linalg.slogdet(o.informationmatrix(pts, w, j) + R)[1] for j in range(5)])
    assert oreg.objective_batch([pts], [w])[0] == pytest.approx(want, rel=1e-13)
    # only the upper triangle of R matters, like every normalized information matrix of the reference
    Ru = np.triu(R) + np.tril(np.full((3, 3), 7.0), -1)
    assert ko.Problem(ko.TPC, [1.0], th).set_regularization(Ru).objective_batch([pts], [w])[0] == \
        oreg.objective_batch([pts], [w])[0]


def test_oracle_vignette_design_needs_regularization():  # docs/src/extend-approximation.md:171-183
    dt = np.stack([dttm(TPC_G1)])
    o = ko.Problem(ko.TPC, [1.0], [TPC_G1], dt=dt)
    assert o.objective_batch([S_TR[0]], [S_TR[1]])[0] == -np.inf
    o.set_regularization(R4)
    assert np.isfinite(o.objective_batch([S_TR[0]], [S_TR[1]])[0])
    gd = o.gateauxderivative(S_TR[0], S_TR[1], np.linspace(0, 48, 4801))
    assert gd.max() < 2e-2     # a 50-iteration PSO result
    # the derivative really is the directional derivative of the regularised objective: with trB = t = tr[A (M + R)]
    # the reference's tr[A (M_z + R)] - trB equals tr[A (M_z - M)].  The regularised matrix is nearly singular, so
    # the curvature is large and the step has to be tiny.
    for z in (1.0, 10.0):
        a = 1e-9
        f1 = o.objective_batch([S_TR[0] + [z]], [list((1 - a) * np.array(S_TR[1])) + [a]])[0]
        f0 = o.objective_batch([S_TR[0] + [z]], [S_TR[1] + [0.0]])[0]
        assert (f1 - f0) / a == pytest.approx(o.gateauxderivative(S_TR[0], S_TR[1], [z])[0], rel=1e-4)


# ---- GPU: the engine ---------------------------------------------------------------------------------------------

def H(model, mp, theta, crit=0, dt=None, pw=None, ub=(1.0,), R=None):
    import kirstine_b200 as kb
    pw = np.full(len(theta), 1.0 / len(theta)) if pw is None else pw
    h = kb.ProblemHandle(model, mp, theta, pw, crit, dt, (0.0,), ub)
    return h.set_regularization(R) if R is not None else h


@pytest.mark.gpu
@pytest.mark.parametrize("crit", [0, 1, 2])
@pytest.mark.parametrize("delta", [False, True])
def test_engine_regularized_tpc_vs_oracle(crit, delta):
    th = wl.tpc_prior(400, seed=9)
    R = spd(3, 2, 1e-3)
    dt = np.stack([dttm(t) for t in th]) if delta else None
    h = H(3, [1.0], th, crit, dt, ub=(48.0,), R=R)
    o = ko.Problem(ko.TPC, [1.0], th, crit=crit, dt=dt).set_regularization(R)
    pts, w = wl.jittered_swarm(30, 4, 0.0, 48.0, seed=4, log_lo=0.05)
    pts[0, 2:], w[0] = pts[0, :2], [0.3, 0.7, 0.0, 0.0]     # a two-point design: singular without R
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)
    assert np.all(np.isfinite(h.objective_batch(pts[:1], w[:1])))
    np.testing.assert_allclose(h.informationmatrix(pts[1], w[1], 7), o.informationmatrix(pts[1], w[1], 7), rtol=RTOL)
    dirs = np.linspace(0, 48, 241)
    for i in (0, 1):
        assert not h.gateaux_prepare(pts[i], w[i])
        want = o.gateauxderivative(pts[i], w[i], dirs)
        np.testing.assert_allclose(h.gateaux_eval(dirs), want, rtol=RTOL, atol=RTOL * np.abs(want).max())
    # switching back to FisherMatrix restores the unregularised numbers
    h.set_regularization(None)
    o.set_regularization(None)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    if crit == 2 and not delta:   # Dexp: exp(-Inf / t) = 0; the oracle's two-point matrix is singular up to rounding only
        assert got[0] == 0.0 and 0.0 <= want[0] < 1e-4
    else:
        assert got[0] == want[0] == -np.inf
    np.testing.assert_allclose(got[1:], want[1:], rtol=RTOL)


@pytest.mark.gpu
def test_engine_regularized_sigemax_uniform_prior():
    # uniform prior weights: the unregularised problem runs the determinant-product kernel, the regularised one
    # must not
    th = wl.sigemax_prior(500, seed=31415)
    R = spd(4, 3, 1e-5)
    h, o = H(2, [1.0], th, R=R), ko.Problem(ko.SIGEMAX, [1.0], th).set_regularization(R)
    pts, w = wl.jittered_swarm(50, 8, 0.0, 1.0, seed=6)
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=RTOL)
    one = ([[[0.4]]], [[1.0]])
    np.testing.assert_allclose(h.objective_batch(*one), o.objective_batch(*one), rtol=RTOL)


@pytest.mark.gpu
def test_engine_regularization_argument_checks():
    import kirstine_b200 as kb
    h = H(3, [1.0], [TPC_G1], ub=(48.0,))
    with pytest.raises(ValueError):
        h.set_regularization(np.eye(4))
    with pytest.raises(kb.EngineError):
        h.set_regularization(np.full((3, 3), np.nan))


@pytest.mark.gpu
def test_engine_solves_vignette_problem():  # docs/src/extend-approximation.md:113-176
    import kirstine_b200 as kb

    def dp_for(trafo, approx):
        return kb.DesignProblem(region=kb.DesignInterval(("time", (0, 48))), criterion=kb.DCriterion(),
                                covariate_parameterization=kb.JustCopy("time"), model=kb.TPCModel(sigma=1),
                                prior_knowledge=kb.PriorSample([kb.TPCParameter(a=4.298, e=0.05884, s=21.80)]),
                                transformation=trafo, normal_approximation=approx)

    dp_tr = dp_for(kb.DeltaMethod(lambda p: dttm(p)), kb.RegularizedFisherMatrix(R4))
    dp_t = dp_for(kb.DeltaMethod(lambda p: dttm(p)), kb.FisherMatrix())
    gold = kb.DesignMeasure([[x] for x in S_TR[0]], S_TR[1])
    assert kb.objective(gold, dp_t) == -np.inf
    assert np.isfinite(kb.objective(gold, dp_tr))
    kb.seed(1357)
    strategy = kb.DirectMaximization(optimizer=kb.Pso(swarmsize=200, iterations=200),
                                     prototype=kb.uniform_design([[0], [24], [48]]))
    s, _ = kb.solve(dp_tr, strategy, maxweight=1e-3, maxdist=1e-2)
    # "efficiency" is not a sample-size ratio under regularisation (vignette :39-52), but the expression still
    # orders designs: ours must be at least as good as the published one
    assert kb.efficiency(s, gold, dp_tr) >= 0.9999
    M = kb.informationmatrix(gold, kb.model(dp_tr), kb.covariate_parameterization(dp_tr), TPC_G1,
                             kb.normal_approximation(dp_tr))
    M0 = kb.informationmatrix(gold, kb.model(dp_t), kb.covariate_parameterization(dp_t), TPC_G1,
                              kb.normal_approximation(dp_t))
    np.testing.assert_allclose(M - M0, R4, atol=1e-12)
    assert np.linalg.matrix_rank(M0) == 2 and np.linalg.matrix_rank(M) == 3
