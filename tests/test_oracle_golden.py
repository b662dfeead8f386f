"""Pins the CPU oracle against the known-answer tests of the reference's own suite.

Every case cites the reference test it restates (paths relative to the reference
checkout).  Runs on CPU (`-m "not gpu"`).
"""
import numpy as np
import pytest

import oracle as ko

TPC_G1 = [4.298, 0.05884, 21.80]
D1_PTS, D1_W = [0.2288, 1.3886, 18.417], [1 / 3, 1 / 3, 1 / 3]
A1_PTS, A1_W = [0.2176, 1.4343, 18.297], [0.2337, 0.3878, 0.3785]


def dauc(p):
    a, e, s = p
    return np.array([[s / a**2, -s / e**2, 1 / e - 1 / a]])


def dttm(p):
    a, e, s = p
    A = a - e
    B = np.log(a / e)
    return np.array([[(A / a - B) / A**2, (B - A / e) / A**2, 0.0]])


def dcmax(p):
    a, e, s = p
    tmax = np.log(a / e) / (a - e)
    A, E = np.exp(-a * tmax), np.exp(-e * tmax)
    F = a * A - e * E
    da_ttm, de_ttm, _ = dttm(p)[0]
    return np.array([[s * (tmax * A - F * da_ttm), s * (-tmax * E + F * de_ttm), E - A]])


def domnibus(p):
    return np.vstack([dauc(p) / np.sqrt(2194), dttm(p) / np.sqrt(0.02815), dcmax(p) / np.sqrt(1.0)])


def tpc(theta, crit=ko.CRIT_D, dt=None, pw=None):
    return ko.Problem(ko.TPC, [1.0], np.atleast_2d(theta), pw, crit=crit, dt=dt)


# ---- test/design-common.jl -------------------------------------------------

def test_tr_prod_upper():  # test/design-common.jl:26-36
    A = np.arange(1, 10, dtype=float).reshape(3, 3).T
    B = np.arange(11, 20, dtype=float).reshape(3, 3).T
    SA = np.triu(A) + np.triu(A, 1).T
    SB = np.triu(B) + np.triu(B, 1).T
    assert ko.tr_prod_upper(A, B) == np.trace(SA @ SB)


def test_informationmatrix_emax():  # test/design-common.jl:53-63
    p = ko.Problem(ko.EMAX, [1.0], [[1, 10, 5]])
    M = p.informationmatrix([0, 5], [0.25, 0.75])
    ref = np.array([[16, 6, -6], [6, 3, -3], [-6, -3, 3]]) / 16
    np.testing.assert_allclose(M, ref, rtol=1e-14)


def test_informationmatrix_vector_unit():  # test/design-common.jl:66-81
    th = TPC_G1
    p = ko.Problem(ko.VU, [0.1, 3], [th], m=3, cp_id=ko.CP_EQUITIME)
    M = p.informationmatrix([5, 20], [0.25, 0.75])
    F = []
    for times, S in (([0, 5, 10], [0.1, 0.6, 1.1]), ([0, 20, 40], [0.1, 2.1, 4.1])):
        jm = p.jacobian(times, th)
        F.append(jm.T @ np.linalg.inv(np.diag(S)) @ jm)
    np.testing.assert_allclose(M, 0.25 * F[0] + 0.75 * F[1], rtol=1e-12)


def test_apply_transformation_values():  # test/design-common.jl:84-106
    M = np.array([[2.0, 1], [1, 3]])
    out, inv = ko.apply_transformation(M, None)
    assert not inv and np.array_equal(np.triu(out), np.triu(M))
    out, inv = ko.apply_transformation(M, np.array([[0.0, 1]]))
    assert inv and np.linalg.inv(out)[0, 0] == pytest.approx(2.5, rel=1e-15)
    out, inv = ko.apply_transformation(M, np.array([[2.0, 0], [-1, 1]]))
    np.testing.assert_allclose(np.linalg.inv(out), [[1.75, 2], [2, 3.0]], rtol=1e-13)


def test_apply_transformation_inverted_input_agrees():  # test/design-common.jl:117-203
    rng = np.random.default_rng(7)
    X = rng.normal(size=(3, 3))
    M = X @ X.T + np.eye(3)
    DT = rng.normal(size=(2, 3))
    a, fa = ko.apply_transformation(M, DT, r_is_inv=False)
    b, fb = ko.apply_transformation(np.linalg.inv(M), DT, r_is_inv=True)
    assert fa and fb
    np.testing.assert_allclose(a, b, rtol=1e-12)
    np.testing.assert_allclose(a, DT @ np.linalg.inv(M) @ DT.T, rtol=1e-12)
    # identity in disguise
    c, _ = ko.apply_transformation(M, np.eye(3))
    np.testing.assert_allclose(np.linalg.inv(c), M, rtol=1e-12)
    # scaling pulls out as a determinant factor
    d, _ = ko.apply_transformation(M, 2 * np.eye(3))
    assert np.linalg.det(d) == pytest.approx(4**3 * np.linalg.det(np.linalg.inv(M)), rel=1e-12)


def test_log_det():  # test/design-common.jl:108-115
    rng = np.random.default_rng(2468)
    A = rng.random((3, 3))
    B = A @ A.T
    assert ko.log_det(B) == pytest.approx(np.log(np.linalg.det(B)), rel=1e-12)
    assert ko.log_det([[1.0, 0.0], [0.0, 0.0]]) == -np.inf


# ---- test/model-nonlinear-regression.jl -------------------------------------

def test_average_fishermatrix_exact():  # test/model-nonlinear-regression.jl:53-83
    th = np.array([1.0, 10, 5])
    p = ko.Problem(ko.EMAX, [1.0], [th])
    M = p.informationmatrix([0, 5], [0.25, 0.75])
    ref = np.zeros((3, 3))
    for w, x in ((0.25, 0.0), (0.75, 5.0)):
        jm = p.jacobian([x], th)
        ref += w * jm.T @ jm
    assert np.array_equal(M, ref)  # exact, like the reference's `==`


def test_vcov_cholesky_is_sqrt():  # test/model-nonlinear-regression.jl:40-51
    # sigma_squared = 36 -> whitening divides rows by 6 -> M scales by 1/36
    th = [1.0, 10, 2]
    a = ko.Problem(ko.EMAX, [36.0], [th]).informationmatrix([1, 2], [0.5, 0.5])
    b = ko.Problem(ko.EMAX, [1.0], [th]).informationmatrix([1, 2], [0.5, 0.5])
    np.testing.assert_allclose(a * 36, b, rtol=1e-14)


# ---- test/criterion-d.jl -----------------------------------------------------

def test_d_functional():  # test/criterion-d.jl:13-28
    mreg, msng = [[1.0, 0.5], [0.5, 2.0]], [[1.0, 0.5], [0.5, 0.25]]
    assert ko.criterion_functional(mreg, False, ko.CRIT_D) == pytest.approx(np.log(1.75), rel=1e-14)
    assert ko.criterion_functional(msng, False, ko.CRIT_D) == -np.inf
    assert ko.criterion_functional(mreg, True, ko.CRIT_D) == pytest.approx(-np.log(1.75), rel=1e-14)
    assert ko.criterion_functional(msng, True, ko.CRIT_D) == -np.inf


def test_d_objective_tpc_locally_optimal():  # test/criterion-d.jl:36-55 (Atkinson et al. 1993)
    p = tpc(TPC_G1)
    assert p.objective([1.0], [1.0]) == -np.inf
    assert p.objective(D1_PTS, D1_W) == pytest.approx(7.3887, abs=1e-5)


def test_d_objective_delta_auc_distributional():  # test/criterion-d.jl:58-77
    # prior needs Julia's RNG stream; the published value is matched distributionally
    rng = np.random.default_rng(4711)
    n = 1000
    th = np.column_stack([4.298 + 2 * 0.5 * (2 * rng.random(n) - 1),
                          0.05884 + 2 * 0.005 * (2 * rng.random(n) - 1), np.full(n, 21.8)])
    dt = np.stack([dauc(t) for t in th])
    p = tpc(th, dt=dt)
    assert p.objective([1.0], [1.0]) == -np.inf
    val = np.exp(-p.objective([0.2449, 1.4950, 18.4903], [0.0129, 0.0387, 0.9484]))
    assert val == pytest.approx(2463.3, rel=3e-2)


def _two_point_prior():
    g1 = np.array(TPC_G1)
    g2 = g1 + np.array([0.5, 0.005, 0.0])
    return g1, g2


def test_d_gateaux_constants_identity():  # test/criterion-d.jl:80-110
    g1, g2 = _two_point_prior()
    p = tpc([g1, g2])
    with pytest.raises(ArithmeticError):
        p.gateaux_constants([1.0122], [1.0])
    A, trB = p.gateaux_constants(D1_PTS, D1_W)
    for j in range(2):
        np.testing.assert_allclose(A[j], np.linalg.inv(p.informationmatrix(D1_PTS, D1_W, j)), rtol=1e-9)
    assert np.all(trB == 3)


def test_d_gateaux_constants_delta():  # test/criterion-d.jl:112-151
    g1, g2 = _two_point_prior()
    J = [dauc(g1), dauc(g2)]
    p = tpc([g1, g2], dt=np.stack(J))
    with pytest.raises(ArithmeticError):
        p.gateaux_constants([1.0122], [1.0])
    A, trB = p.gateaux_constants(D1_PTS, D1_W)
    for j in range(2):
        iM = np.linalg.inv(p.informationmatrix(D1_PTS, D1_W, j))
        B = J[j].T @ np.linalg.inv(J[j] @ iM @ J[j].T) @ J[j]
        np.testing.assert_allclose(A[j], iM @ B @ iM, rtol=1e-8)
    assert np.all(trB == 1)


def test_d_gateauxderivative_equivalence_theorem():  # test/criterion-d.jl:153-195
    dirs = np.linspace(0, 48, 21)
    pi = tpc(TPC_G1)
    pd = tpc(TPC_G1, dt=np.eye(3)[None])
    assert np.all(np.isnan(pi.gateauxderivative([1.0], [1.0], dirs)))
    assert np.all(np.isnan(pd.gateauxderivative([1.0], [1.0], dirs)))
    gi = pi.gateauxderivative(D1_PTS, D1_W, dirs)
    assert gi.max() <= 0
    assert np.all(np.abs(pi.gateauxderivative(D1_PTS, D1_W, D1_PTS)) < 1e-4)
    np.testing.assert_allclose(gi, pd.gateauxderivative(D1_PTS, D1_W, dirs), rtol=1e-9, atol=1e-12)


# ---- test/criterion-a.jl -----------------------------------------------------

def test_a_functional():  # test/criterion-a.jl:12-27
    mreg, msng = [[1.0, 0.5], [0.5, 2.0]], [[1.0, 0.5], [0.5, 0.25]]
    assert ko.criterion_functional(mreg, False, ko.CRIT_A) == pytest.approx(-12 / 7, rel=1e-14)
    assert ko.criterion_functional(mreg, True, ko.CRIT_A) == pytest.approx(-3, rel=1e-15)
    with pytest.raises(ArithmeticError):
        ko.criterion_functional(msng, False, ko.CRIT_A)


def test_a_objective_identity_equals_delta_identity():  # test/criterion-a.jl:37-62
    a = tpc(TPC_G1, crit=ko.CRIT_A).objective(A1_PTS, A1_W)
    b = tpc(TPC_G1, crit=ko.CRIT_A, dt=np.eye(3)[None]).objective(A1_PTS, A1_W)
    assert a == pytest.approx(b, rel=1e-10)


def test_a_objective_omnibus():  # test/criterion-a.jl:65-83 (Atkinson et al. 1993)
    p = tpc(TPC_G1, crit=ko.CRIT_A, dt=domnibus(TPC_G1)[None])
    assert p.objective([1.0], [1.0]) == -np.inf
    assert -p.objective(A1_PTS, A1_W) == pytest.approx(6.883, abs=1e-3)


def test_a_gateaux_constants_identity():  # test/criterion-a.jl:86-117
    g1, g2 = _two_point_prior()
    p = tpc([g1, g2], crit=ko.CRIT_A)
    with pytest.raises(ArithmeticError):
        p.gateaux_constants([1.0122], [1.0])
    A, trB = p.gateaux_constants(D1_PTS, D1_W)
    for j in range(2):
        iM = np.linalg.inv(p.informationmatrix(D1_PTS, D1_W, j))
        np.testing.assert_allclose(A[j], iM @ iM, rtol=1e-8)
        assert trB[j] == pytest.approx(np.trace(iM), rel=1e-9)


def test_a_gateaux_constants_delta():  # test/criterion-a.jl:119-155
    g1, g2 = _two_point_prior()
    J = [dauc(g1), dauc(g2)]
    p = tpc([g1, g2], crit=ko.CRIT_A, dt=np.stack(J))
    A, trB = p.gateaux_constants(D1_PTS, D1_W)
    for j in range(2):
        iM = np.linalg.inv(p.informationmatrix(D1_PTS, D1_W, j))
        np.testing.assert_allclose(A[j], iM @ J[j].T @ J[j] @ iM, rtol=1e-8)
        assert trB[j] == pytest.approx(np.trace(J[j] @ iM @ J[j].T), rel=1e-9)


def test_a_gateauxderivative():  # test/criterion-a.jl:157-216
    dirs = np.linspace(0, 48, 21)
    pi = tpc(TPC_G1, crit=ko.CRIT_A)
    pd = tpc(TPC_G1, crit=ko.CRIT_A, dt=np.eye(3)[None])
    assert np.all(np.isnan(pi.gateauxderivative([1.0], [1.0], dirs)))
    np.testing.assert_allclose(pi.gateauxderivative(D1_PTS, D1_W, dirs),
                               pd.gateauxderivative(D1_PTS, D1_W, dirs), rtol=1e-8, atol=1e-10)
    po = tpc(TPC_G1, crit=ko.CRIT_A, dt=domnibus(TPC_G1)[None])
    assert np.all(np.isnan(po.gateauxderivative([1.0], [1.0], dirs)))
    assert po.gateauxderivative(A1_PTS, A1_W, dirs).max() <= 0
    assert np.all(np.abs(po.gateauxderivative(A1_PTS, A1_W, A1_PTS)) < 1e-2)


# ---- analytic Emax solution (test/example-emax.jl:49-61) -----------------------

def test_emax_locally_optimal_gateaux():
    # x* = 2.5 on [0,10] with ec50 = 5, uniform weights: the Gateaux derivative must vanish
    # at the support and be non-positive elsewhere (equivalence theorem).
    p = ko.Problem(ko.EMAX, [1.0], [[1, 10, 5]])
    g = p.gateauxderivative([0, 2.5, 10], [1 / 3] * 3, np.linspace(0, 10, 101))
    assert g.max() < 1e-12
    assert np.all(np.abs(p.gateauxderivative([0, 2.5, 10], [1 / 3] * 3, [0, 2.5, 10])) < 1e-12)


# ---- independent numpy restatement (cross-check of the C arithmetic) ----------

def _np_objective(jac, pts, w, thetas, pw, sigma2=1.0, crit="D", dts=None):
    acc = 0.0
    for j, th in enumerate(thetas):
        M = sum(wk * np.outer(jac(x, th), jac(x, th)) / sigma2 for x, wk in zip(pts, w))
        if dts is not None:
            MT = np.linalg.inv(dts[j] @ np.linalg.inv(M) @ dts[j].T)
        else:
            MT = M
        acc += pw[j] * (np.linalg.slogdet(MT)[1] if crit == "D" else -np.trace(np.linalg.inv(MT)))
    return acc


def _sigemax_jac(x, th):
    e0, emax, ed50, h = th
    xp, ep = x**h, ed50**h
    A = xp / (xp + ep)
    B = ep * emax / (xp + ep)
    return np.array([1.0, A, -A * B * h / ed50, 0.0 if x == 0 else A * B * np.log(x / ed50)])


def test_sigemax_objective_vs_numpy():  # docs/src/tutorial.md:127-142,190-196
    rng = np.random.default_rng(31415)
    th = np.array([1, 2, 0.4, 5]) + np.array([0.5, 0.5, 0.05, 0.5]) * rng.normal(size=(200, 4))
    pw = np.full(200, 1 / 200)
    pts = np.linspace(0, 1, 8)
    w = np.full(8, 1 / 8)
    p = ko.Problem(ko.SIGEMAX, [1.0], th)
    assert p.objective(pts, w) == pytest.approx(_np_objective(_sigemax_jac, pts, w, th, pw), rel=1e-11)
    # A criterion with DeltaMethod -> ED90 (docs/src/autodiff.md:257-266)
    rho = 90.0
    a = (100 / rho - 1) ** (1 / th[:, 3])
    dt = np.zeros((200, 1, 4))
    dt[:, 0, 2] = 1 / a
    dt[:, 0, 3] = th[:, 2] * np.log(100 / rho - 1) / (th[:, 3] ** 2 * a)
    pa = ko.Problem(ko.SIGEMAX, [1.0], th, crit=ko.CRIT_A, dt=dt)
    assert pa.objective(pts, w) == pytest.approx(
        _np_objective(_sigemax_jac, pts, w, th, pw, crit="A", dts=dt), rel=1e-10)


def test_objective_batch_threads_match_serial():
    rng = np.random.default_rng(3)
    th = np.column_stack([4.298 + (2 * rng.random(50) - 1), 0.05884 + 0.01 * (2 * rng.random(50) - 1),
                          np.full(50, 21.8)])
    p = tpc(th)
    pts = rng.random((6, 6, 1)) * 48
    w = rng.dirichlet(np.ones(6), size=6)
    a = p.objective_batch(pts, w, nthreads=1)
    b = p.objective_batch(pts, w, nthreads=4)
    assert np.array_equal(a, b)
    assert a[0] == p.objective(pts[0], w[0])
    # shards of the prior sum to the whole (multi-GPU partition)
    parts = p.objective_range(pts[0], w[0], 0, 20) + p.objective_range(pts[0], w[0], 20, 50)
    assert parts == pytest.approx(a[0], rel=1e-13)
