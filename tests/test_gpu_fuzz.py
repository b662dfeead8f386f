"""Randomised differential test: shapes the hand-written cases do not enumerate -- every swarm size around the
particle-tile and warp boundaries, every row count (so every remainder / group combination of the row loop), zero
weights and trivial rows in random places, odd draw counts, non-uniform prior weights, all criteria and
transformation sizes -- engine against oracle, arbitrated by the 80-bit referee where the oracle itself is only
good to the bar."""
import numpy as np
import pytest

import oracle as ko
from tests import referee as rf
from tests import workloads as wl

pytestmark = pytest.mark.gpu
RTOL = 1e-10

FAMILIES = {   # name: (model id, model params, prior, ub, log_lo, r)
    "emax": (1, [1.3], wl.emax_prior, 10.0, None, 3),
    "sigemax": (2, [0.8], wl.sigemax_prior, 1.0, None, 4),
    "tpc": (3, [1.1], wl.tpc_prior, 48.0, 0.05, 3),
}


def _orthonormal_rows(n, t, r, seed):
    rng = np.random.default_rng(seed)
    q = np.linalg.qr(rng.normal(size=(n, r, r)))[0]
    return np.ascontiguousarray(np.transpose(q, (0, 2, 1))[:, :t, :])


import os  # noqa: E402

@pytest.mark.parametrize("seed", range(int(os.environ.get("KST_FUZZ_CASES", "24"))))
def test_random_shapes_against_oracle(seed):
    import kirstine_b200 as kb
    rng = np.random.default_rng(1000 + seed)
    name = list(FAMILIES)[seed % 3]
    model, mp, prior, ub, log_lo, r = FAMILIES[name]
    N = int(rng.choice([1, 2, 31, 33, 255, 257, 700]))
    S = int(rng.choice([1, 2, 7, 8, 9, 15, 17, 33]))
    K = int(rng.integers(r, r + 9))                       # r .. r+8 rows: all remainders of the 2- and 4-row groups
    crit = int(rng.integers(0, 3))
    t = int(rng.integers(0, r + 1))                       # 0 = Identity, else DeltaMethod with t rows
    th = prior(N, seed=seed + 5)
    pw = rng.dirichlet(np.ones(N)) if rng.random() < 0.5 else np.full(N, 1.0 / N)
    dt = _orthonormal_rows(N, t, r, seed) if t else None
    pts, w = wl.jittered_swarm(S, K, 0.0, ub, seed=seed, log_lo=log_lo)
    # sprinkle zero weights (dropped rows) and, for the families that have them, trivial rows (dose / time 0);
    # keep at least r live, distinct rows so that most particles stay regular
    for s in range(S):
        extra = K - r
        nz = int(rng.integers(0, extra + 1))
        idx = rng.choice(K, size=nz, replace=False)
        w[s, idx] = 0.0
        w[s] /= w[s].sum()
        if extra - nz > 0 and rng.random() < 0.5:
            live = np.flatnonzero(w[s] > 0)
            pts[s, rng.choice(live), 0] = 0.0
    h = kb.ProblemHandle(model, mp, th, pw, crit, dt, (0.0,), (ub,))
    o = ko.Problem(model, mp, th, pw, crit=crit, dt=dt)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    fin = np.isfinite(want) & np.isfinite(got)
    assert fin.mean() >= 0.5, (name, N, S, K, crit, t)
    # every particle goes through assert_parity: Inf / NaN classes must agree (unless the scaled condition number is
    # beyond 1e12, where the side of "singular" a double lands on is rounding luck), the 1e-10 bar holds outright for
    # cond < 1e6 (against the oracle or, where the oracle's own rounding is the larger error, the 80-bit referee), and
    # beyond that the engine may be at most 8 x as far from the referee as the reference arithmetic itself
    cond = {}

    def cond_fn(i):
        if i not in cond:
            cond[i] = rf.cond_max(model, mp, pts[i], w[i], th, dt=dt)
        return cond[i]

    rf.assert_parity(got, want, lambda i: rf.objective(model, mp, pts[i], w[i], th, pw, crit=crit, dt=dt), rtol=RTOL,
                     cond_fn=cond_fn)
    fin &= np.array([cond_fn(i) < 1e6 if fin[i] else False for i in range(S)])
    if not fin.any():
        return
    # Gateaux derivative at the first regular particle
    s0 = int(np.flatnonzero(fin)[0])
    dirs = np.linspace(ub * 0.01, ub, 37)
    if not h.gateaux_prepare(pts[s0], w[s0]):
        gd, wg = h.gateaux_eval(dirs), o.gateauxderivative(pts[s0], w[s0], dirs)
        rf.assert_parity(gd, wg, lambda i: rf.gateaux(model, mp, pts[s0], w[s0], [dirs[i]], th, pw, crit=crit, dt=dt)[0],
                         rtol=RTOL, scale=np.abs(wg).max())


@pytest.mark.parametrize("seed", range(int(os.environ.get("KST_FUZZ_CASES", "24")) // 2))
def test_random_shapes_vector_units(seed):
    """The same for the families with several rows per design point: VUMod (m = 2..5 rows at times j * delta, variance
    growing with time) and the dose-time-response model with equidistant times on a two-dimensional region."""
    import kirstine_b200 as kb
    rng = np.random.default_rng(5000 + seed)
    vu = seed % 2 == 0
    N = int(rng.choice([1, 3, 32, 65, 300]))
    S = int(rng.choice([1, 3, 8, 9, 16, 19]))
    m = int(rng.integers(2, 6))
    crit = int(rng.integers(0, 3))
    pw = rng.dirichlet(np.ones(N)) if rng.random() < 0.5 else np.full(N, 1.0 / N)
    if vu:
        model, mp, r, cp, cpp, lb, ub = 4, [0.1], 3, 1, [], [0.0], [24.0]
        th = wl.tpc_prior(N, seed=seed + 2)
        K = int(rng.integers(2, 7))
        pts, w = wl.jittered_swarm(S, K, 0.0, 24.0, seed=seed, log_lo=0.2)
        omp, cov = [0.1, m], None
    else:
        model, mp, r, cp, cpp, lb, ub = 5, [1.0], 5, 5, [], [0.0, 0.0], [100.0, 6.0]
        th = wl.dtr_prior(N, seed=seed + 2)
        K = int(rng.integers(3, 8))
        pts, w = wl.jittered_swarm_nd(S, K, lb, ub, seed=seed)
        omp, cov = [1.0], (5, [])
    t = int(rng.integers(0, r + 1))
    dt = _orthonormal_rows(N, t, r, seed + 9) if t else None
    for s_ in range(S):       # zero weights in random places, at least two points left
        nz = int(rng.integers(0, max(1, K - 2)))
        w[s_, rng.choice(K, size=nz, replace=False)] = 0.0
        w[s_] /= w[s_].sum()
    h = kb.ProblemHandle(model, mp, th, pw, crit, dt, lb, ub, m=m, cp_id=cp, cp_params=cpp)
    o = ko.Problem(model, omp, th, pw, crit=crit, dt=dt, m=m, d=len(lb), cp_id=cp, cp_params=cpp)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    rf.assert_parity(got, want, lambda i: rf.objective(model, mp, pts[i], w[i], th, pw, crit=crit, dt=dt, m=m, cov=cov),
                     rtol=RTOL, cond_fn=lambda i: rf.cond_max(model, mp, pts[i], w[i], th, m, cov, dt=dt))


@pytest.mark.parametrize("seed", range(int(os.environ.get("KST_FUZZ_CASES", "24")) // 2))
def test_random_shapes_extension_seams(seed):
    """Logistic families, the regularised normal approximation and the Dexp criterion on random shapes (no 80-bit
    referee for these: well-separated designs, engine against oracle at the 1e-10 bar, Gateaux grid included)."""
    import kirstine_b200 as kb
    rng = np.random.default_rng(9000 + seed)
    kind = seed % 3
    N = int(rng.choice([1, 5, 33, 129, 400]))
    S = int(rng.choice([1, 2, 8, 11, 24]))
    crit = int(rng.integers(0, 3))
    pw = rng.dirichlet(np.ones(N)) if rng.random() < 0.5 else np.full(N, 1.0 / N)
    reg = None
    if kind == 0:      # LinLogReg, d = 1..3
        d = int(rng.integers(1, 4))
        model, mp, r = 6, [0.0], d + 1
        th = np.column_stack([rng.normal(-1.0, 0.5, N)] + [rng.normal(1.5, 0.7, N) for _ in range(d)])
        lb, ub = [0.0] * d, [1.0] * d
        K = int(rng.integers(r + 1, r + 6))
        pts = rng.uniform(0, 1, (S, K, d))
        w = rng.dirichlet(np.full(K, 4.0), S)
        o = ko.Problem(ko.LINLOGREG, [], th, pw, crit=crit, d=d)
    elif kind == 1:    # SinLogReg
        model, mp, r, d = 7, [0.0], 3, 1
        th = np.column_stack([rng.uniform(0.5, 1.5, N), rng.uniform(1.0, 2.0, N), rng.uniform(0.0, 0.5, N)])
        lb, ub = [0.0], [1.0]
        K = int(rng.integers(4, 9))
        pts, w = wl.jittered_swarm(S, K, 0.0, 1.0, seed=seed)
        o = ko.Problem(ko.SINLOGREG, [], th, pw, crit=crit, d=1)
    else:              # regularised sigmoid Emax / TPC
        sig = rng.random() < 0.5
        model, mp, r, d = (2, [0.9], 4, 1) if sig else (3, [1.2], 3, 1)
        th = wl.sigemax_prior(N, seed=seed) if sig else wl.tpc_prior(N, seed=seed)
        lb, ub = [0.0], [1.0 if sig else 48.0]
        K = int(rng.integers(2, 8))                 # fewer points than parameters are fine with R
        pts, w = wl.jittered_swarm(S, K, 0.0, ub[0], seed=seed, log_lo=None if sig else 0.05)
        q = rng.normal(size=(r, r))
        reg = 10.0 ** rng.uniform(-4, -1) * (q @ q.T + r * np.eye(r))
        o = ko.Problem(model, mp, th, pw, crit=crit).set_regularization(reg)
    h = kb.ProblemHandle(model, mp, th, pw, crit, None, lb, ub)
    if reg is not None:
        h.set_regularization(reg)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    np.testing.assert_allclose(got[fin], want[fin], rtol=1e-9)
    if fin.any():
        s0 = int(np.flatnonzero(fin)[0])
        dirs = rng.uniform(np.array(lb) + 0.01, ub, (29, len(lb)))
        if not h.gateaux_prepare(pts[s0], w[s0]):
            wg = o.gateauxderivative(pts[s0], w[s0], dirs)
            np.testing.assert_allclose(h.gateaux_eval(dirs), wg, rtol=1e-8, atol=1e-9 * np.abs(wg).max())
