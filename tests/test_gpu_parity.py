"""Parity of the CUDA engine (through the C ABI) with the CPU oracle.

Tolerance: BASELINE.json north_star asks for <= 1e-10 relative error in FP64 on objective,
Gateaux-derivative and information-matrix values; RTOL below is that bar.
"""
import numpy as np
import pytest

import oracle as ko
from tests import referee as rf
from tests import workloads as wl

pytestmark = pytest.mark.gpu
RTOL = 1e-10

kb = None


@pytest.fixture(scope="module", autouse=True)
def _engine():
    global kb
    import kirstine_b200
    kb = kirstine_b200
    assert kb.device_count() >= 1, "no CUDA device: GPU tests must run on the B200 box"
    yield


def H(model, mp, theta, crit=0, dt=None, pw=None, lb=(0.0,), ub=(1.0,), m=1, cp=0):
    pw = np.full(len(theta), 1.0 / len(theta)) if pw is None else pw
    return kb.ProblemHandle(model, mp, theta, pw, crit, dt, lb, ub, m=m, cp_id=cp)


def O(model, mp, theta, crit=0, dt=None, pw=None, m=1, cp=0):
    return ko.Problem(model, mp, theta, pw, crit=crit, dt=dt, m=m, cp_id=cp)


def close(a, b, rtol=RTOL):
    a, b = np.asarray(a, float), np.asarray(b, float)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    assert np.array_equal(np.isinf(a), np.isinf(b))
    fin = np.isfinite(a)
    np.testing.assert_allclose(a[fin], b[fin], rtol=rtol, atol=0)
    assert np.array_equal(a[~fin & ~np.isnan(a)], b[~fin & ~np.isnan(b)])


TPC_G1 = [4.298, 0.05884, 21.80]
D1_PTS, D1_W = [0.2288, 1.3886, 18.417], [1 / 3, 1 / 3, 1 / 3]
A1_PTS, A1_W = [0.2176, 1.4343, 18.297], [0.2337, 0.3878, 0.3785]


# ---- information matrix golden values (test/design-common.jl:53-81) ---------------------------

def test_informationmatrix_emax_golden():
    h = H(1, [1.0], [[1, 10, 5]], ub=(10.0,))
    M = h.informationmatrix([0, 5], [0.25, 0.75])
    np.testing.assert_allclose(M, np.array([[16, 6, -6], [6, 3, -3], [-6, -3, 3]]) / 16, rtol=1e-14)


def test_informationmatrix_vector_unit_golden():
    th = [TPC_G1]
    M = H(4, [0.1], th, ub=(24.0,), m=3, cp=1).informationmatrix([5, 20], [0.25, 0.75])
    close(M, O(ko.VU, [0.1, 3], th, m=3, cp=ko.CP_EQUITIME).informationmatrix([5, 20], [0.25, 0.75]), 1e-12)


@pytest.mark.parametrize("model,mp,theta,ub", [
    (1, [2.5], wl.emax_prior(7), 10.0), (2, [0.7], wl.sigemax_prior(7), 1.0), (3, [1.3], wl.tpc_prior(7), 48.0)])
def test_informationmatrix_vs_oracle(model, mp, theta, ub):
    h, o = H(model, mp, theta, ub=(ub,)), O(model, mp, theta)
    pts, w = wl.random_swarm(3, 5, 0.0, ub, seed=11)
    for j in (0, 3, 6):
        close(h.informationmatrix(pts[1], w[1], j), o.informationmatrix(pts[1], w[1], j))


# ---- objective: known answers, then the five BASELINE configs at oracle-sized N ----------------

def test_objective_tpc_known_answers():
    h = H(3, [1.0], [TPC_G1], ub=(48.0,))
    assert h.objective_batch([[[1.0]]], [[1.0]])[0] == -np.inf   # one_point_design([1]): singular
    fx = h.objective_batch([np.reshape(D1_PTS, (3, 1))], [D1_W])
    assert fx[0] == pytest.approx(7.3887, abs=1e-5)      # test/criterion-d.jl:36-55
    dom = np.array([__import__("tests.test_oracle_golden", fromlist=["x"]).domnibus(TPC_G1)])
    ha = H(3, [1.0], [TPC_G1], crit=1, dt=dom, ub=(48.0,))
    fa = ha.objective_batch([np.reshape(A1_PTS, (3, 1))], [A1_W])
    assert -fa[0] == pytest.approx(6.883, abs=1e-3)      # test/criterion-a.jl:65-83
    assert ha.objective_batch([[[1.0]]], [[1.0]])[0] == -np.inf


def _cfg(cfg, n):
    if cfg == "1":   # Emax, D, Identity, K = 3 on [0, 10]
        return dict(model=1, mp=[1.0], theta=wl.emax_prior(n), ub=10.0, K=3)
    if cfg == "1b":  # tutorial: sigmoid Emax, D, Identity, K = 8 on [0, 1]
        return dict(model=2, mp=[1.0], theta=wl.sigemax_prior(n), ub=1.0, K=8)
    if cfg == "2":   # sigmoid Emax, A, DeltaMethod -> ED90, K = 6
        th = wl.sigemax_prior(n, seed=2)
        return dict(model=2, mp=[1.0], theta=th, ub=1.0, K=6, crit=1, dt=wl.ded90_jacobians(th))
    if cfg == "3":   # TPC, D, Identity, K = 6 on [0, 48]
        return dict(model=3, mp=[1.0], theta=wl.tpc_prior(n), ub=48.0, K=6)
    if cfg == "4":   # vector unit m = 3, D, Identity, region [0, 24]
        return dict(model=4, mp=[0.1], theta=wl.tpc_prior(n, seed=4), ub=24.0, K=3, m=3, cp=1)
    raise KeyError(cfg)


def _swarm(c, S, seed):
    log_lo = 0.05 if c["model"] == 3 else None
    return wl.jittered_swarm(S, c["K"], 0.0, c["ub"], seed=seed, log_lo=log_lo)


def _parity(c, h, o, pts, w, crit, dt, m, pw=None):
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w, nthreads=8)
    pw = np.full(len(c["theta"]), 1.0 / len(c["theta"])) if pw is None else pw
    rf.assert_parity(got, want, lambda i: rf.objective(c["model"], c["mp"], pts[i], w[i], c["theta"], pw, crit=crit,
                                                       dt=dt, m=m), rtol=RTOL,
                     cond_fn=lambda i: rf.cond_max(c["model"], c["mp"], pts[i], w[i], c["theta"], m, dt=dt))
    return got, want


@pytest.mark.parametrize("cfg,n,S", [("1", 1000, 50), ("1b", 1000, 40), ("2", 2000, 32), ("3", 3000, 24),
                                     ("4", 1000, 24)])
def test_objective_batch_vs_oracle(cfg, n, S):
    c = _cfg(cfg, n)
    m, cp, crit, dt = c.get("m", 1), c.get("cp", 0), c.get("crit", 0), c.get("dt")
    h = H(c["model"], c["mp"], c["theta"], crit=crit, dt=dt, ub=(c["ub"],), m=m, cp=cp)
    omp = c["mp"] + ([m] if c["model"] == 4 else [])
    o = O(c["model"], omp, c["theta"], crit=crit, dt=dt, m=m, cp=cp)
    pts, w = _swarm(c, S, 5)
    # boundary hits as the swarm produces them: exact lower / upper bounds
    pts[2, 0] = 0.0
    pts[3, -1] = c["ub"]
    got, want = _parity(c, h, o, pts, w, crit, dt, m)
    # these well-separated designs are all well conditioned (scaled cond(M) < 1e6), so assert_parity held every one
    # of them to the north-star bar outright: within 1e-10 of the oracle, or of the 80-bit referee where the oracle's
    # own rounding is the larger error -- no slack rule, no skipped element
    conds = np.array([rf.cond_max(c["model"], c["mp"], pts[i], w[i], c["theta"], m, dt=dt) for i in range(S)])
    assert conds.max() < 1e6 and rf.last_report["slack"] == 0 and rf.last_report["singular_luck"] == 0, rf.last_report
    assert np.isfinite(want).all() and np.isfinite(got).all()
    # the reference's own random initialiser (i.i.d. points): the same rule wherever cond < 1e6, the referee with a
    # slack of 8 x the oracle's own error beyond that, Inf / NaN classes must agree unless cond >= 1e12
    pts, w = wl.random_swarm(S, c["K"] + 2, 0.0, c["ub"], seed=6)
    _parity(c, h, o, pts, w, crit, dt, m)


def test_objective_zero_weight_point_is_dropped():
    # K = 4 sigmoid-Emax design with one zero weight leaves 3 points for 4 parameters: singular by construction
    th = wl.sigemax_prior(50)
    pts, w = wl.jittered_swarm(3, 4, 0.0, 1.0, seed=1)
    w[1, 2] = 0.0
    w[1] /= w[1].sum()
    got = H(2, [1.0], th).objective_batch(pts, w)
    assert got[1] == -np.inf and np.isfinite(got[0]) and np.isfinite(got[2])
    # with K = 5 the zero-weight point just vanishes
    pts, w = wl.jittered_swarm(2, 5, 0.0, 1.0, seed=2)
    w[0, 4] = 0.0
    w[0] /= w[0].sum()
    a = H(2, [1.0], th).objective_batch(pts, w)[0]
    b = H(2, [1.0], th).objective_batch(pts[:1, :4], w[:1, :4])[0]
    assert a == pytest.approx(b, rel=1e-13)


def _orthonormal_rows(n, t, r, seed):
    rng = np.random.default_rng(seed)
    q = np.linalg.qr(rng.normal(size=(n, r, r)))[0]
    return np.ascontiguousarray(np.transpose(q, (0, 2, 1))[:, :t, :]) * rng.uniform(0.5, 2.0, size=(n, t, 1))


@pytest.mark.parametrize("crit", [0, 1])
@pytest.mark.parametrize("t", [0, 1, 2, 3])
def test_objective_all_criterion_transformation_variants(crit, t):
    th = wl.tpc_prior(257, seed=9)
    dt = None if t == 0 else _orthonormal_rows(len(th), t, 3, t)
    c = dict(model=3, mp=[0.8], theta=th, ub=48.0, K=4)
    h, o = H(3, [0.8], th, crit=crit, dt=dt, ub=(48.0,)), O(3, [0.8], th, crit=crit, dt=dt)
    pts, w = _swarm(c, 9, 3)
    got, want = _parity(c, h, o, pts, w, crit, dt, 1)
    assert np.all(np.isfinite(got))


@pytest.mark.parametrize("crit", [0, 1])
@pytest.mark.parametrize("t", [1, 2, 3, 4])
def test_objective_sigemax_delta_variants(crit, t):
    th = wl.sigemax_prior(300, seed=t)
    dt = _orthonormal_rows(len(th), t, 4, 10 + t)
    c = dict(model=2, mp=[1.3], theta=th, ub=1.0, K=6)
    h, o = H(2, [1.3], th, crit=crit, dt=dt), O(2, [1.3], th, crit=crit, dt=dt)
    pts, w = _swarm(c, 10, 4)
    got, _ = _parity(c, h, o, pts, w, crit, dt, 1)
    assert np.all(np.isfinite(got))


def test_objective_a_identity_equals_delta_identity():  # test/criterion-a.jl:37-62
    a = H(3, [1.0], [TPC_G1], crit=1, ub=(48.0,)).objective_batch([np.reshape(A1_PTS, (3, 1))], [A1_W])
    b = H(3, [1.0], [TPC_G1], crit=1, dt=np.eye(3)[None], ub=(48.0,)).objective_batch(
        [np.reshape(A1_PTS, (3, 1))], [A1_W])
    assert a[0] == pytest.approx(b[0], rel=1e-10)


def test_objective_singular_and_nan_semantics():
    th = wl.sigemax_prior(64)
    pw = np.full(64, 1 / 63.0)
    pw[5] = 0.0
    # particle 0 has 3 rows for 4 parameters (the dose-0 row and two more): singular by construction
    pts = np.array([[[0.0], [0.0], [0.3], [0.6]], [[0.0], [0.3], [0.6], [1.0]]])
    w = np.full((2, 4), 0.25)
    for crit in (0, 1):
        h, o = H(2, [1.0], th, crit=crit, pw=pw), O(2, [1.0], th, crit=crit, pw=pw)
        got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
        close(got[1:], want[1:])
        assert np.isfinite(got[1]) and not np.isfinite(got[0])
    # D / Identity: singular design with a zero prior weight gives 0 * -Inf = NaN, like the reference
    assert np.isnan(H(2, [1.0], th, pw=pw).objective_batch(pts, w)[0])
    # A / Identity and DeltaMethod: the exception path returns -Inf regardless of weights
    assert H(2, [1.0], th, crit=1, pw=pw).objective_batch(pts, w)[0] == -np.inf
    assert H(2, [1.0], th, dt=np.tile(np.eye(4)[None], (64, 1, 1)), pw=pw).objective_batch(pts, w)[0] == -np.inf


def test_objective_ragged_sizes():
    # draws not a multiple of the CTA width, swarm not a multiple of the particle tile, K = 1
    th = wl.emax_prior(1)
    close(H(1, [1.0], th, ub=(10.0,)).objective_batch([[[3.0]]], [[1.0]]),
          O(1, [1.0], th).objective_batch([[[3.0]]], [[1.0]]))
    th = wl.emax_prior(777)
    pts, w = wl.jittered_swarm(13, 5, 0.0, 10.0, seed=2)
    c = dict(model=1, mp=[4.0], theta=th)
    _parity(c, H(1, [4.0], th, ub=(10.0,)), O(1, [4.0], th), pts, w, 0, None, 1)


def test_objective_reproducible_bitwise():
    th = wl.sigemax_prior(5000)
    h = H(2, [1.0], th)
    pts, w = wl.jittered_swarm(33, 8, 0.0, 1.0, seed=8)
    a, b = h.objective_batch(pts, w), h.objective_batch(pts, w)
    assert np.array_equal(a, b)
    # permuting the swarm permutes the result exactly
    perm = np.random.default_rng(0).permutation(33)
    assert np.array_equal(h.objective_batch(pts[perm], w[perm]), a[perm])


def test_objective_full_size_properties():
    """BASELINE config-5 sized prior (10^6 draws): size-independent properties, and the oracle on a few particles."""
    th = wl.sigemax_prior(1_000_000, seed=5)
    pts, w = wl.jittered_swarm(16, 8, 0.0, 1.0, seed=1)
    full = H(2, [1.0], th).objective_batch(pts, w)
    # linearity in the prior: two half shards with weights 1/N sum to the whole
    half = len(th) // 2
    pw = np.full(len(th), 1.0 / len(th))
    a = H(2, [1.0], th[:half], pw=pw[:half]).objective_batch(pts, w)
    b = H(2, [1.0], th[half:], pw=pw[half:]).objective_batch(pts, w)
    np.testing.assert_allclose(a + b, full, rtol=1e-12)
    # sigma only shifts the D objective by -r log sigma^2
    s = H(2, [2.0], th).objective_batch(pts, w)
    np.testing.assert_allclose(s, full - 4 * np.log(4.0), rtol=1e-12)
    # the oracle itself on the full prior (all host threads; a second of CPU time for these 16 particles)
    import os
    want = O(2, [1.0], th).objective_batch(pts, w, nthreads=os.cpu_count() or 1)
    np.testing.assert_allclose(full, want, rtol=RTOL)
    # a bounded sample of the same prior agrees with the oracle
    sub = th[:4000]
    _parity(dict(model=2, mp=[1.0], theta=sub), H(2, [1.0], sub), O(2, [1.0], sub), pts, w, 0, None, 1)


# ---- Gateaux derivative ---------------------------------------------------------------------------

@pytest.mark.parametrize("crit", [0, 1])
@pytest.mark.parametrize("delta", [False, True])
def test_gateaux_constants_vs_oracle(crit, delta):  # test/criterion-{d,a}.jl gateaux_constants
    g1 = np.array(TPC_G1)
    th = np.array([g1, g1 + [0.5, 0.005, 0.0]])
    dauc = __import__("tests.test_oracle_golden", fromlist=["x"]).dauc
    dt = np.stack([dauc(t) for t in th]) if delta else None
    h, o = H(3, [1.0], th, crit=crit, dt=dt, ub=(48.0,)), O(3, [1.0], th, crit=crit, dt=dt)
    assert h.gateaux_prepare([1.0122], [1.0]) is True       # singular design
    assert h.gateaux_prepare(D1_PTS, D1_W) is False
    A, trB = h.gateaux_constants()
    Ao, trBo = o.gateaux_constants(D1_PTS, D1_W)
    # both draws are well conditioned at this design: the north-star bar, entrywise against the matrix's largest entry
    # for the near-zero entries
    np.testing.assert_allclose(A, Ao, rtol=RTOL, atol=RTOL * np.abs(Ao).max(axis=(1, 2), keepdims=True).min())
    np.testing.assert_allclose(trB, trBo, rtol=RTOL)


@pytest.mark.parametrize("cfg,n", [("1", 500), ("1b", 500), ("3", 800), ("4", 1000)])
@pytest.mark.parametrize("crit", [0, 1])
def test_gateaux_eval_vs_oracle(cfg, n, crit):
    c = _cfg(cfg, n)
    m, cp = c.get("m", 1), c.get("cp", 0)
    h = H(c["model"], c["mp"], c["theta"], crit=crit, ub=(c["ub"],), m=m, cp=cp)
    omp = c["mp"] + ([m] if c["model"] == 4 else [])
    o = O(c["model"], omp, c["theta"], crit=crit, m=m, cp=cp)
    at_p = np.linspace(0, c["ub"], 5)
    at_w = np.array([0.1, 0.2, 0.3, 0.25, 0.15])
    dirs = np.linspace(0, c["ub"], 301)
    assert not h.gateaux_prepare(at_p, at_w)
    got = h.gateaux_eval(dirs)
    want = o.gateauxderivative(at_p, at_w, dirs, nthreads=8)
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=RTOL * np.abs(want).max())


def test_gateaux_delta_variants_vs_oracle():
    th = wl.tpc_prior(300, seed=12)
    pw = np.full(len(th), 1.0 / len(th))
    dirs = np.exp(np.linspace(np.log(0.05), np.log(48), 61))
    for crit in (0, 1):
        for t in (1, 2, 3):
            dt = _orthonormal_rows(len(th), t, 3, 20 + t)
            h, o = H(3, [1.0], th, crit=crit, dt=dt, ub=(48.0,)), O(3, [1.0], th, crit=crit, dt=dt)
            assert not h.gateaux_prepare(D1_PTS, D1_W)
            got, want = h.gateaux_eval(dirs), o.gateauxderivative(D1_PTS, D1_W, dirs)
            rf.assert_parity(got, want, lambda i: rf.gateaux(3, [1.0], D1_PTS, D1_W, [dirs[i]], th, pw, crit=crit,
                                                             dt=dt)[0], rtol=RTOL, scale=np.abs(want).max())


def test_gateaux_equivalence_theorem_and_nan_fill():  # test/criterion-d.jl:153-195
    h = H(3, [1.0], [TPC_G1], ub=(48.0,))
    dirs = np.linspace(0, 48, 21)
    assert h.gateaux_prepare([1.0], [1.0])
    assert np.all(np.isnan(h.gateaux_eval(dirs)))
    assert not h.gateaux_prepare(D1_PTS, D1_W)
    g = h.gateaux_eval(dirs)
    assert g.max() <= 0
    assert np.all(np.abs(h.gateaux_eval(D1_PTS)) < 1e-4)
    hd = H(3, [1.0], [TPC_G1], dt=np.eye(3)[None], ub=(48.0,))
    hd.gateaux_prepare(D1_PTS, D1_W)
    np.testing.assert_allclose(hd.gateaux_eval(dirs), g, rtol=1e-9, atol=1e-12)


def test_gateaux_grid_config4_size():
    """config 4: 10^4-direction grid x 1000 draws, vector unit m = 3; full grid vs a strided oracle sample."""
    c = _cfg("4", 1000)
    h = H(c["model"], c["mp"], c["theta"], ub=(24.0,), m=3, cp=1)
    o = O(ko.VU, [0.1, 3], c["theta"], m=3, cp=1)
    at_p, at_w = [2.0, 8.0, 20.0], [0.3, 0.3, 0.4]
    dirs = np.linspace(0, 24, 10_000)
    assert not h.gateaux_prepare(at_p, at_w)
    got = h.gateaux_eval(dirs)
    want = o.gateauxderivative(at_p, at_w, dirs[::50], nthreads=8)
    np.testing.assert_allclose(got[::50], want, rtol=RTOL, atol=RTOL * np.abs(want).max())


def test_gateaux_grid_large_prior():
    """The same grid against 10^5 draws (draw range split over the grid, chunks prefetched through registers): a
    strided sample of directions against the oracle, and the whole grid against two half priors (linearity)."""
    import os
    th = wl.tpc_prior(100_000, seed=4)
    pw = np.full(len(th), 1.0 / len(th))
    h = H(4, [0.1], th, ub=(24.0,), m=3, cp=1)
    o = O(ko.VU, [0.1, 3], th, m=3, cp=1)
    at_p, at_w = [2.0, 8.0, 20.0], [0.3, 0.3, 0.4]
    dirs = np.linspace(0, 24, 10_000)
    assert not h.gateaux_prepare(at_p, at_w)
    got = h.gateaux_eval(dirs)
    want = o.gateauxderivative(at_p, at_w, dirs[::250], nthreads=os.cpu_count() or 1)
    np.testing.assert_allclose(got[::250], want, rtol=RTOL, atol=RTOL * np.abs(want).max())
    half = len(th) // 2
    parts = []
    for sl in (slice(0, half), slice(half, None)):
        hh = H(4, [0.1], th[sl], pw=pw[sl], ub=(24.0,), m=3, cp=1)
        assert not hh.gateaux_prepare(at_p, at_w)
        parts.append(hh.gateaux_eval(dirs))
    np.testing.assert_allclose(parts[0] + parts[1], got, rtol=1e-11, atol=1e-11 * np.abs(got).max())


# ---- error behaviour of the boundary ----------------------------------------------------------------

def test_abi_errors():
    with pytest.raises(kb.EngineError, match="unregistered model"):
        H(99, [1.0], [[1, 2, 3]])
    with pytest.raises(kb.EngineError, match="parameter dimension"):
        H(1, [1.0], [[1, 2, 3, 4]])
    with pytest.raises(kb.EngineError, match="upper bounds"):
        H(1, [1.0], [[1, 2, 3]], lb=(1.0,), ub=(1.0,))
    h = H(1, [1.0], [[1, 10, 5]], ub=(10.0,))
    with pytest.raises(kb.EngineError, match="before gateaux_prepare"):
        h.gateaux_eval([1.0])
    with pytest.raises(kb.EngineError, match="at least 2"):
        h.pso_init(1, 3, [[0.0], [5.0], [10.0]], [1 / 3] * 3)
    with pytest.raises(kb.EngineError, match="c1 \\+ c2 > 4"):
        h.pso_init(5, 3, [[0.0], [5.0], [10.0]], [1 / 3] * 3, c1=1, c2=1)
    with pytest.raises(kb.EngineError, match="outside design region"):
        h.pso_init(5, 3, [[0.0], [5.0], [11.0]], [1 / 3] * 3)


def test_engine_math_accuracy():
    """kst_exp / kst_rcp / kst_log (csrc/kst_math.cuh) against mpmath over the ranges the kernels use."""
    import mpmath as mp
    mp.mp.prec = 200
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-708, 708, 4000), rng.uniform(-30, 30, 4000), rng.normal(0, 1e-3, 500),
                        [0.0, -0.0, 707.9, -707.9, 720.0, -720.0, 5000.0, -5000.0, np.nan]])
    e, _, _ = kb._ffi.math_selftest(x)
    worst = 0.0
    for xi, ei in zip(x[:8500:7], e[:8500:7]):
        ref = mp.exp(mp.mpf(float(xi)))
        worst = max(worst, float(abs(mp.mpf(float(ei)) - ref) / ref) / 2.0 ** -52)
    assert worst <= 1.0, worst
    assert e[-9] == 1.0 and e[-8] == 1.0
    assert e[-7] == pytest.approx(np.exp(707.9), rel=1e-15) and e[-6] == pytest.approx(np.exp(-707.9), rel=1e-15)
    # beyond |x| = 708 the result saturates at a finite, normal value; NaN stays NaN
    assert np.isfinite(e[-5]) and e[-5] > 1e300 and 0 < e[-4] < 1e-300
    assert np.isfinite(e[-3]) and e[-3] > 1e300 and 0 < e[-2] < 1e-300 and np.isnan(e[-1])
    y = np.concatenate([np.exp(rng.uniform(-600, 600, 4000)), -np.exp(rng.uniform(-5, 5, 100)), 1 + rng.random(1000)])
    _, r, l = kb._ffi.math_selftest(y)
    assert np.max(np.abs(r - 1.0 / y) / np.abs(1.0 / y)) <= 2.0 * 2.0 ** -52
    worst_abs, worst_rel = 0.0, 0.0
    for yi, li in zip(y[::9], l[::9]):
        ref = mp.log(mp.mpf(abs(float(yi))))
        err = float(abs(mp.mpf(float(li)) - ref))
        worst_abs = max(worst_abs, err / max(1.0, abs(float(ref))))
    assert worst_abs <= 3e-16, worst_abs
    near1 = 1.0 + rng.uniform(-1e-3, 1e-3, 500)
    _, _, l1 = kb._ffi.math_selftest(near1)
    assert np.max(np.abs(l1 - np.log(near1))) <= 3e-16


def test_objective_exponent_range_paths():
    """The D/Identity/uniform-weights kernel keeps two copies of its draw loop: particles that can prove their exp
    arguments stay within +-700 for every draw skip the exponent clamp, the others keep it.  One swarm mixes
    both kinds (doses down to 1e-250, where h log(ed50/x) overflows the exponent range) and a prior with one
    steep draw flips more particles to the clamped copy; results must not depend on the path."""
    th = wl.sigemax_prior(400, seed=8)
    pts, w = wl.jittered_swarm(16, 8, 0.0, 1.0, seed=8)
    pts[1, 0], pts[2, 3], pts[3, 5], pts[4, 1] = 1e-250, 1e-120, 1e-30, 1e-300
    h, o = H(2, [1.0], th), O(2, [1.0], th)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    pw = np.full(len(th), 1.0 / len(th))
    rf.assert_parity(got, want, lambda i: rf.objective(2, [1.0], pts[i], w[i], th, pw), rtol=RTOL,
                     cond_fn=lambda i: rf.cond_max(2, [1.0], pts[i], w[i], th))
    # one draw with a steep Hill coefficient pushes |h log(ed50 / x)| past 700 for the small doses of every
    # particle: all of them take the clamped copy now
    th2 = th.copy()
    th2[7, 3] = 12.0
    got2 = H(2, [1.0], th2).objective_batch(pts, w)
    want2 = O(2, [1.0], th2).objective_batch(pts, w)
    rf.assert_parity(got2, want2, lambda i: rf.objective(2, [1.0], pts[i], w[i], th2, pw), rtol=RTOL,
                     cond_fn=lambda i: rf.cond_max(2, [1.0], pts[i], w[i], th2))
    # two-exponential family: a prior draw with a huge elimination rate sends exp(-a t) far below the range
    tt = wl.tpc_prior(300, seed=2)
    tt[5, 0] = 60.0
    ptt, wt = wl.jittered_swarm(12, 5, 0.0, 48.0, seed=3, log_lo=0.05)
    close(H(3, [1.0], tt, ub=(48.0,)).objective_batch(ptt, wt), O(3, [1.0], tt).objective_batch(ptt, wt), 1e-9)


def test_objective_swarm_size_limit():
    """More particle tiles than gridDim.y can hold is refused with KST_ELIMIT, not launched."""
    h = H(1, [1.0], wl.emax_prior(4), ub=(10.0,))
    S = 65535 * 8 + 1
    with pytest.raises(kb.EngineError, match="at most"):
        h.objective_batch(np.full((S, 1, 1), 1.0), np.ones((S, 1)))
    big = h.objective_batch(np.full((65535 * 8, 1, 1), 1.0), np.ones((65535 * 8, 1)))     # the limit itself runs
    assert np.all(big == -np.inf)                                                        # one point, three parameters
