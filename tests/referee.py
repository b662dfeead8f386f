"""Extended-precision referee (x87 80-bit long double, 64-bit mantissa) for objective and
Gateaux-derivative values.

Used to arbitrate engine-vs-oracle disagreements on ill-conditioned designs: random swarm
particles often have cond(M) >= 1e8, where the reference's own double arithmetic is only good to
1e-8 and a 1e-10 engine-vs-oracle comparison is meaningless.  `assert_parity` accepts a value when
it is within the north-star tolerance of the oracle, or at least as close to the referee as the
oracle is (up to a small factor); where the design is well conditioned (cond < 1e6) the bar holds outright, and
Inf / NaN results must agree in class.  Test infrastructure only.
"""
import numpy as np

LD = np.longdouble


def dtr_covariate(cp, cpp, z, m):
    """(dose, [time_1..time_m]) of the DTR covariate parameterizations (docs/src/dtr.md)."""
    z = np.atleast_1d(np.asarray(z, dtype=np.float64))
    if cp == 3:
        return cpp[0], [z[0]] + [0.0] * (m - 1)
    if cp == 4:
        return z[0], list(cpp[:m])
    if cp == 5:
        return z[0], [float(j) * z[1] for j in range(m)]
    if cp == 6:
        t = [0.0, z[1]]
        for _ in range(2, m):
            t.append(cpp[0] * t[-1])
        return z[0], t[:m]
    return z[0], [z[1]] + [0.0] * (m - 1)


def rows(model, mp, z, theta, m=1, cov=None):
    """Whitened Jacobian rows of one design point for all draws: (N, m, r) long double."""
    th = theta.astype(LD)
    N = len(th)
    if model == 5:  # DTR, mp = sigma; cov = (cp_id, cp_params)
        dose, times = dtr_covariate(cov[0], cov[1], z, m)
        a, e, e0, emax, ec50 = th.T
        dose = LD(dose)
        out = []
        for t in times:
            t = LD(t)
            A, B = np.exp(-a * t), np.exp(-e * t)
            C, rd = B - A, e - a
            den = dose * C * a - rd * ec50
            g = np.stack([-dose * ec50 * emax * (A * t * a * rd + e * C) / den ** 2,
                          dose * a * ec50 * emax * (B * t * rd + C) / den ** 2, np.ones(N, LD), dose * a * C / den,
                          dose * a * emax * C * rd / den ** 2], axis=1)
            out.append(g / abs(LD(mp[0])))
        return np.stack(out, axis=1)
    z = LD(np.atleast_1d(z)[0])
    if model == 1:  # Emax, mp = sigma_squared
        e0, emax, ec50 = th.T
        g = np.stack([np.ones(N, LD), z / (z + ec50), -emax * z / (z + ec50) ** 2], axis=1)[:, None, :]
        return g / np.sqrt(LD(mp[0]))
    if model == 2:  # sigmoid Emax, mp = sigma
        e0, emax, ed50, h = th.T
        if z == 0:
            g = np.stack([np.ones(N, LD), np.zeros(N, LD), np.zeros(N, LD), np.zeros(N, LD)], axis=1)
        else:
            xp, ep = z ** h, ed50 ** h
            A = xp / (xp + ep)
            B = ep * emax / (xp + ep)
            g = np.stack([np.ones(N, LD), A, -A * B * h / ed50, A * B * np.log(z / ed50)], axis=1)
        return g[:, None, :] / abs(LD(mp[0]))
    if model in (3, 4):  # TPC (mp = sigma) / VU (mp = sigma_squared), times (i) * z
        a, e, s = th.T
        out = []
        for i in range(m):
            t = z if model == 3 else LD(i) * z
            A, E = np.exp(-a * t), np.exp(-e * t)
            g = np.stack([A * s * t, -E * s * t, E - A], axis=1)
            var = LD(mp[0]) ** 2 if model == 3 else LD(mp[0]) * (1 + t)
            out.append(g / np.sqrt(var))
        return np.stack(out, axis=1)
    if model == 8:  # VUMod with AR(1) errors, mp = (sigma_squared, rho): rows whitened with the dense unit covariance
        a, e, s = th.T
        t = np.array([LD(i) * z for i in range(m)], dtype=LD)
        J = np.stack([np.stack([np.exp(-a * ti) * s * ti, -np.exp(-e * ti) * s * ti, np.exp(-e * ti) - np.exp(-a * ti)],
                               axis=1) for ti in t], axis=1)                       # (N, m, 3)
        S = np.array([[LD(mp[0]) * LD(mp[1]) ** abs(i - j) * np.sqrt((1 + t[i]) * (1 + t[j])) for j in range(m)]
                      for i in range(m)], dtype=LD)
        U = np.zeros((m, m), dtype=LD)                                             # S = U'U
        for j in range(m):
            U[j, j] = np.sqrt(S[j, j] - np.sum(U[:j, j] ** 2))
            for i in range(j + 1, m):
                U[j, i] = (S[j, i] - np.sum(U[:j, j] * U[:j, i])) / U[j, j]
        X = np.zeros_like(J)                                                       # U' X = J
        for i in range(m):
            X[:, i, :] = (J[:, i, :] - np.einsum("l,nlr->nr", U[:i, i], X[:, :i, :])) / U[i, i]
        return X
    raise KeyError(model)


def infomatrix(model, mp, pts, w, theta, m=1, cov=None):
    w = np.asarray(w, dtype=np.float64).ravel()
    pts = np.asarray(pts, dtype=np.float64).reshape(len(w), -1)
    M = 0
    for z, wk in zip(pts, w):
        g = rows(model, mp, z, theta, m, cov)
        M = M + LD(wk) * np.einsum("nia,nib->nab", g, g)
    return M


def inv_spd(M):
    """Batched inverse by Gauss-Jordan (no pivoting; M is SPD) and a flag for non-positive pivots."""
    M = M.copy()
    n = M.shape[-1]
    I = np.broadcast_to(np.eye(n, dtype=LD), M.shape).copy()
    ok = np.ones(M.shape[0], bool)
    logdet = np.zeros(M.shape[0], LD)
    for c in range(n):
        piv = M[:, c, c].copy()
        ok &= piv > 0
        with np.errstate(all="ignore"):
            logdet += np.log(np.where(piv > 0, piv, 1))
            M[:, c, :] /= piv[:, None]
            I[:, c, :] /= piv[:, None]
            for a in range(n):
                if a != c:
                    f = M[:, a, c].copy()
                    M[:, a, :] -= f[:, None] * M[:, c, :]
                    I[:, a, :] -= f[:, None] * I[:, c, :]
    return I, ok, logdet


def objective(model, mp, pts, w, theta, pw, crit=0, dt=None, m=1, cov=None):
    M = infomatrix(model, mp, pts, w, theta, m, cov)
    iM, ok, logdet = inv_spd(M)
    if not ok.all():
        return -np.inf
    if dt is None:
        val = logdet if crit == 0 else -np.trace(iM, axis1=1, axis2=2)
    else:
        D = dt.astype(LD)
        C = np.einsum("nai,nij,nbj->nab", D, iM, D)
        if crit == 1:
            val = -np.trace(C, axis1=1, axis2=2)
        else:
            _, ok2, ld2 = inv_spd(C)
            if not ok2.all():
                return -np.inf
            val = -ld2
    return float(np.sum(pw.astype(LD) * val))


def gateaux(model, mp, at_pts, at_w, dirs, theta, pw, crit=0, dt=None, m=1, cov=None):
    M = infomatrix(model, mp, at_pts, at_w, theta, m, cov)
    iM, ok, _ = inv_spd(M)
    assert ok.all()
    r = M.shape[-1]
    if dt is None:
        if crit == 0:
            A, trB = iM, np.full(len(M), LD(r))
        else:
            A, trB = np.einsum("nab,nbc->nac", iM, iM), np.trace(iM, axis1=1, axis2=2)
    else:
        D = dt.astype(LD)
        C = np.einsum("nai,nij,nbj->nab", D, iM, D)
        if crit == 0:
            W, okc, _ = inv_spd(C)
            assert okc.all()
            A = np.einsum("nij,naj,nab,nbk,nkl->nil", iM, D, W, D, iM)
            trB = np.full(len(M), LD(D.shape[1]))
        else:
            A = np.einsum("nij,naj,nak,nkl->nil", iM, D, D, iM)
            trB = np.trace(C, axis1=1, axis2=2)
    out = []
    d = np.asarray(at_pts, dtype=np.float64).reshape(len(np.ravel(at_w)), -1).shape[1]
    for z in np.asarray(dirs, dtype=np.float64).reshape(-1, d):
        Md = infomatrix(model, mp, [z], [1.0], theta, m, cov)
        out.append(float(np.sum(pw.astype(LD) * (np.einsum("nab,nab->n", A, Md) - trB))))
    return np.array(out)


def _cond_scaled(M):
    """2-norm condition numbers of a batch of symmetric matrices after scaling to unit diagonal; Inf where a matrix
    is not finite or has a non-positive diagonal entry."""
    with np.errstate(all="ignore"):
        dg = np.sqrt(np.einsum("nii->ni", M))
        Ms = np.asarray(M / (dg[:, :, None] * dg[:, None, :]), dtype=np.float64)
    out = np.full(len(Ms), np.inf)
    good = np.all(np.isfinite(Ms), axis=(1, 2))
    if good.any():
        with np.errstate(all="ignore"):
            c = np.linalg.cond(Ms[good])
        out[good] = np.where(np.isfinite(c), c, np.inf)
    return out


def cond_max(model, mp, pts, w, theta, m=1, cov=None, sample=4096, dt=None):
    """Largest 2-norm condition number of the design's information matrix over a strided sample of the prior draws,
    from the extended-precision matrices, after scaling to unit diagonal (Cholesky-based log det / inverse are
    invariant under diagonal scaling, so this is the number their accuracy depends on; the raw matrices of the
    compartment models carry column scales like exp(-a t) that say nothing about the design).  With a DeltaMethod
    transformation (dt: (N, t, r) Jacobians) the criterion is taken on DT M^-1 DT', whose own scaled condition number
    counts as well: a dense DT mixes the widely different scales of M^-1 and the t x t matrix can be close to rank
    one although M is harmless.  Inf when a matrix is not finite or has a non-positive diagonal entry.
    The sample covers every draw of the test-sized priors (<= 4096): one ill-conditioned draw can carry the whole
    error of a prior-averaged A criterion (found by the fuzz test at a draw a 48-draw sample had missed)."""
    th = np.asarray(theta, dtype=np.float64)
    step = max(1, len(th) // sample)
    th = th[::step]
    with np.errstate(all="ignore"):
        M = infomatrix(model, mp, pts, w, th, m, cov)
    c = float(np.max(_cond_scaled(M)))
    if dt is not None and np.isfinite(c):
        iM, ok, _ = inv_spd(M)
        if not ok.all():
            return np.inf
        D = np.asarray(dt)[::step].astype(LD)
        with np.errstate(all="ignore"):
            C = np.einsum("nai,nij,nbj->nab", D, iM, D)
        c = max(c, float(np.max(_cond_scaled(C))))
    return c


def _same_class(g, o):
    if np.isnan(o) or np.isnan(g):
        return np.isnan(o) and np.isnan(g)
    return g == o  # both +Inf or both -Inf


# counters of the last assert_parity call, for the callers' messages / reports
last_report = {}


def assert_parity(got, oracle, referee_fn, rtol=1e-10, slack=8.0, scale=None, cond_fn=None, cond_good=1e6,
                  cond_singular=1e12):
    """got/oracle: arrays.  referee_fn(i) -> extended-precision value of element i (called only where got and oracle
    disagree beyond rtol).  scale: magnitude for absolute comparison (Gateaux values cross zero), default |oracle|
    elementwise.  cond_fn(i) -> condition number of element i's design (cond_max), or None.

    Rules, per element:
      * engine within rtol of the oracle                                              -> ok;
      * oracle or engine not finite: the other must be the same Inf / NaN class, unless cond_fn says the design is
        numerically singular (cond >= cond_singular), where which side of "singular" double rounding lands on is
        luck in both implementations; without cond_fn the classes must match;
      * otherwise the 80-bit referee decides: engine within rtol of the referee       -> ok;
        cond_fn given and cond < cond_good: nothing else is accepted (the north-star bar holds outright);
        else the design is ill-conditioned for double arithmetic: the oracle must itself be off the bar
        (>= 0.1 bar from the referee) and the engine no worse than `slack` times the oracle's own error, or -- the
        oracle's error being one lucky or unlucky realisation of rounding noise -- within the forward error bound
        cond * 2^-53 (relative) that no double-precision factorisation can beat.
    The counts of each outcome are left in `last_report`."""
    got, oracle = np.asarray(got, float), np.asarray(oracle, float)
    assert got.shape == oracle.shape
    sc = np.abs(oracle) if scale is None else np.full(got.shape, float(scale))
    bad = []
    rep = {"n": int(got.size), "direct": 0, "class_match": 0, "singular_luck": 0, "referee": 0, "slack": 0, "cond_bound": 0}
    for i in range(got.size):
        g, o = got.flat[i], oracle.flat[i]
        if not (np.isfinite(o) and np.isfinite(g)):
            if _same_class(g, o):
                rep["class_match"] += 1
            elif cond_fn is not None and cond_fn(i) >= cond_singular:
                rep["singular_luck"] += 1
            else:
                bad.append((i, g, o, "class"))
            continue
        if abs(g - o) <= rtol * sc.flat[i]:
            rep["direct"] += 1
            continue
        ref = referee_fn(i)
        c = cond_fn(i) if cond_fn is not None else None
        if not np.isfinite(ref):
            # the 80-bit arithmetic itself finds the design singular while both doubles are finite
            if c is not None and c >= cond_singular:
                rep["singular_luck"] += 1
            else:
                bad.append((i, g, o, ref))
            continue
        eg, eo = abs(g - ref), abs(o - ref)
        bar = rtol * max(sc.flat[i], abs(ref))
        if eg <= bar:
            rep["referee"] += 1
            continue
        if c is not None and c < cond_good:
            bad.append((i, g, o, ref, c))
            continue
        if eo >= 0.1 * bar and eg <= slack * eo:
            rep["slack"] += 1
            continue
        # the oracle's error is one realisation of rounding noise and can be small by luck: a backward-stable
        # factorisation in double cannot promise better than cond * 2^-53 relative, so that much is accepted as well
        if c is not None and eg <= c * 2.0 ** -53 * max(sc.flat[i], abs(ref), 1.0):
            rep["cond_bound"] += 1
            continue
        bad.append((i, g, o, ref) + ((c,) if c is not None else ()))
    last_report.clear()
    last_report.update(rep)
    assert not bad, "engine off the bar (index, engine, oracle, referee[, cond]): %r; outcomes %r" % (bad[:5], rep)
