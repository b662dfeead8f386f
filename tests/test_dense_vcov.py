"""General unit covariance: update_model_workspace! factors an arbitrary m x m Sigma(c_k) with
potrf! and average_fishermatrix! whitens every unit's Jacobian with trsm! (src/model-nonlinear-regression.jl:39-49,
60-86).  None of the reference's own models has a non-diagonal Sigma, so the registered example is VUMod
(test/example-vector.jl) with AR(1)-correlated errors: oracle (CPU, its dpotf2 + dtrsm loops) against closed forms
here, engine against the oracle in the GPU tests below."""
import numpy as np
import pytest

import oracle as ko
from tests import referee as rf
from tests import workloads as wl

RTOL = 1e-10


def _sigma(sig2, rho, t):
    m = len(t)
    return np.array([[sig2 * rho ** abs(i - j) * np.sqrt((1 + t[i]) * (1 + t[j])) for j in range(m)] for i in range(m)])


def _jac(theta, t):
    a, e, s = theta
    return np.array([[np.exp(-a * ti) * s * ti, -np.exp(-e * ti) * s * ti, np.exp(-e * ti) - np.exp(-a * ti)] for ti in t])


@pytest.mark.parametrize("m", [3, 13])
def test_oracle_dense_covariance_against_closed_form(m):
    th = wl.tpc_prior(5, seed=1)
    o = ko.Problem(ko.VUAR, [0.1, 0.6], th, m=m, cp_id=ko.CP_EQUITIME)
    pts, w = np.array([[0.4], [1.5], [3.0]]), np.array([0.2, 0.5, 0.3])
    for j in (0, 4):
        want = sum(wk * _jac(th[j], np.arange(m) * z[0]).T @ np.linalg.solve(_sigma(0.1, 0.6, np.arange(m) * z[0]),
                                                                              _jac(th[j], np.arange(m) * z[0]))
                   for z, wk in zip(pts, w))
        np.testing.assert_allclose(o.informationmatrix(pts, w, j), want, rtol=1e-9)
        ref = np.asarray(rf.infomatrix(8, [0.1, 0.6], pts, w, th[j:j + 1], m=m)[0], dtype=float)
        np.testing.assert_allclose(o.informationmatrix(pts, w, j), ref, rtol=1e-10)


def test_oracle_rho_zero_is_vumod():
    th = wl.tpc_prior(40, seed=2)
    pts, w = wl.jittered_swarm(5, 4, 0.0, 6.0, seed=3, log_lo=0.2)
    for crit in (0, 1):
        a = ko.Problem(ko.VUAR, [0.1, 0.0], th, crit=crit, m=3, cp_id=ko.CP_EQUITIME).objective_batch(pts, w)
        b = ko.Problem(ko.VU, [0.1, 3], th, crit=crit, m=3, cp_id=ko.CP_EQUITIME).objective_batch(pts, w)
        np.testing.assert_allclose(a, b, rtol=1e-13)


# ---- engine ------------------------------------------------------------------------------------------------------

def _pair(th, rho, m, crit=0, dt=None, pw=None, ub=6.0):
    import kirstine_b200 as kb
    pw = np.full(len(th), 1.0 / len(th)) if pw is None else pw
    h = kb.ProblemHandle(kb._ffi.MODEL_VUAR, [0.1, rho], th, pw, crit, dt, [0.0], [ub], m=m, cp_id=kb._ffi.CP_EQUITIME)
    o = ko.Problem(ko.VUAR, [0.1, rho], th, pw, crit=crit, dt=dt, m=m, cp_id=ko.CP_EQUITIME)
    return h, o


@pytest.mark.gpu
@pytest.mark.parametrize("m,K,ub", [(3, 4, 6.0), (13, 5, 1.0), (1, 5, 6.0), (16, 8, 0.8)])
@pytest.mark.parametrize("crit", [0, 1])
def test_engine_dense_covariance_objective_and_matrix(m, K, ub, crit):
    th = wl.tpc_prior(300, seed=7)
    pw = np.random.default_rng(m).dirichlet(np.ones(len(th))) if crit else None
    h, o = _pair(th, 0.55, m, crit, pw=pw, ub=ub)
    pts, w = wl.jittered_swarm(19, K, 0.0, ub, seed=4, log_lo=0.03 * ub)
    w[2, 1] = 0.0                                  # a dropped unit
    w[2] /= w[2].sum()
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w, nthreads=8)
    pwv = np.full(len(th), 1.0 / len(th)) if pw is None else pw
    rf.assert_parity(got, want, lambda i: rf.objective(8, [0.1, 0.55], pts[i], w[i], th, pwv, crit=crit, m=m), rtol=RTOL,
                     cond_fn=lambda i: rf.cond_max(8, [0.1, 0.55], pts[i], w[i], th, m))
    assert np.isfinite(got).all() or m == 1
    for j in (0, 123):
        M = h.informationmatrix(pts[1], w[1], j)
        np.testing.assert_allclose(M, o.informationmatrix(pts[1], w[1], j), rtol=1e-10, atol=1e-13 * np.abs(M).max())


@pytest.mark.gpu
def test_engine_rho_zero_is_the_diagonal_family():
    import kirstine_b200 as kb
    th = wl.tpc_prior(500, seed=9)
    pw = np.full(len(th), 1.0 / len(th))
    pts, w = wl.jittered_swarm(17, 4, 0.0, 6.0, seed=1, log_lo=0.2)
    dense = kb.ProblemHandle(kb._ffi.MODEL_VUAR, [0.1, 0.0], th, pw, 0, None, [0.0], [6.0], m=3, cp_id=1)
    diag = kb.ProblemHandle(kb._ffi.MODEL_VU, [0.1], th, pw, 0, None, [0.0], [6.0], m=3, cp_id=1)
    np.testing.assert_allclose(dense.objective_batch(pts, w), diag.objective_batch(pts, w), rtol=1e-12)
    with pytest.raises(kb.EngineError, match="rho"):
        kb.ProblemHandle(kb._ffi.MODEL_VUAR, [0.1, 1.0], th, pw, 0, None, [0.0], [6.0], m=3, cp_id=1)
    # too few rows for three parameters: singular by construction
    one = kb.ProblemHandle(kb._ffi.MODEL_VUAR, [0.1, 0.3], th, pw, 0, None, [0.0], [6.0], m=2, cp_id=1)
    assert one.objective_batch([[[1.0]]], [[1.0]])[0] == -np.inf


@pytest.mark.gpu
@pytest.mark.parametrize("crit,t", [(0, 0), (1, 0), (0, 2), (1, 1)])
def test_engine_dense_covariance_gateaux_and_delta(crit, t):
    th = wl.tpc_prior(200, seed=11)
    dt = None
    if t:
        q = np.linalg.qr(np.random.default_rng(t).normal(size=(len(th), 3, 3)))[0]
        dt = np.ascontiguousarray(np.transpose(q, (0, 2, 1))[:, :t, :])
    h, o = _pair(th, -0.4, 3, crit, dt)
    pts, w = wl.jittered_swarm(9, 4, 0.0, 6.0, seed=2, log_lo=0.2)
    np.testing.assert_allclose(h.objective_batch(pts, w), o.objective_batch(pts, w), rtol=1e-9)
    at_p, at_w = pts[0], w[0]
    dirs = np.exp(np.linspace(np.log(0.05), np.log(6.0), 301))
    assert not h.gateaux_prepare(at_p, at_w)
    A, trB = h.gateaux_constants()
    Ao, trBo = o.gateaux_constants(at_p, at_w)
    np.testing.assert_allclose(A, Ao, rtol=1e-8, atol=1e-10 * np.abs(Ao).max())
    np.testing.assert_allclose(trB, trBo, rtol=1e-9)
    got, want = h.gateaux_eval(dirs), o.gateauxderivative(at_p, at_w, dirs, nthreads=8)
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-9 * np.abs(want).max())


@pytest.mark.gpu
def test_engine_dense_covariance_solve_and_equivalence_theorem():
    import kirstine_b200 as kb
    prior = kb.PriorSample(wl.tpc_prior(200, seed=5))
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(delta=(0, 6)), model=kb.VUModAR1(0.1, 0.5, 4),
                          covariate_parameterization=kb.EquiTime(), prior_knowledge=prior)
    kb.seed(11)
    d, r = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(iterations=120, swarmsize=60),
                                              prototype=kb.uniform_design([[0.3], [1.0], [3.0]])),
                    maxweight=1e-3, maxdist=1e-3)
    res = kb.optimization_result(r)
    assert np.all(np.diff(res.trace_fx) >= 0) and np.isfinite(res.maximum)
    gd = kb.gateauxderivative(d, [kb.one_point_design([x]) for x in np.linspace(0.01, 6, 300)], dp)
    assert gd.max() < 5e-3
    # both ways through the swarm give the same run
    import os
    kb.seed(11)
    os.environ["KST_NO_FUSED"] = "1"
    try:
        d2, _ = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(iterations=120, swarmsize=60),
                                                   prototype=kb.uniform_design([[0.3], [1.0], [3.0]])),
                         maxweight=1e-3, maxdist=1e-3)
    finally:
        del os.environ["KST_NO_FUSED"]
    assert d == d2
