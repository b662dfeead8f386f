"""Host-side design measures and design regions.

Mirrors the data types of the reference that sit either side of the hot path
(src/designmeasure.jl, src/designregion.jl).  These are O(K^2) host utilities that run once per
solve; the particle format the GPU sees is `DesignMeasure.points` ((K, d) row-major here, which is
byte-identical to Julia's d x K column-major matrix) plus `DesignMeasure.weights`.
"""
import math

import numpy as np


class DesignInterval:
    """DesignInterval{N} (src/designregion.jl:22-51): a box in R^N.

    DesignInterval(("dose", (0, 1)), ("time", (0, 24))) or DesignInterval(dose=(0, 1)).
    """

    def __init__(self, *name_bounds, **kw):
        pairs = list(name_bounds) + list(kw.items())
        if len(pairs) == 3 and not isinstance(pairs[0], tuple):
            names, lb, ub = pairs  # (names, lowerbounds, upperbounds)
        else:
            names = [p[0] for p in pairs]
            lb = [p[1][0] for p in pairs]
            ub = [p[1][1] for p in pairs]
        if not (len(names) == len(lb) == len(ub)):
            raise ValueError("lengths of name and bounds must be identical")
        self.name = tuple(names)
        self.lowerbound = tuple(float(x) for x in lb)
        self.upperbound = tuple(float(x) for x in ub)
        if any(u <= l for l, u in zip(self.lowerbound, self.upperbound)):
            raise ValueError("upper bounds must be strictly larger than lower bounds")

    @property
    def N(self):
        return len(self.name)

    def __repr__(self):
        return "DesignInterval(%r, %r, %r)" % (self.name, self.lowerbound, self.upperbound)


def dimension(dr):
    return dr.N


def lowerbound(dr):
    return dr.lowerbound


def upperbound(dr):
    return dr.upperbound


def dimnames(dr):
    return dr.name


def isinside(dp, dr):  # src/designregion.jl:99-101
    dp = np.asarray(dp, dtype=float)
    return bool(np.all(dp >= dr.lowerbound) and np.all(dp <= dr.upperbound))


class DesignMeasure:
    """DesignMeasure (src/designmeasure.jl:19-42): K design points in R^d with weights.

    DesignMeasure(points, weights) with points (K, d) or a list of K points, or
    DesignMeasure(([1], 0.2), ([42], 0.3), ([9], 0.5)) from (designpoint, weight) pairs.
    """

    def __init__(self, *args):
        if len(args) == 2 and not (isinstance(args[0], tuple) and len(args[0]) == 2 and np.isscalar(args[0][1])):
            pts, w = args
        else:
            pts = [p for p, _ in args]
            w = [x for _, x in args]
        lens = {len(np.atleast_1d(p)) for p in pts}
        if len(lens) > 1:
            raise ValueError("design points must have identical lengths")
        pts = np.array([np.atleast_1d(np.asarray(p, dtype=np.float64)) for p in pts], dtype=np.float64)
        if pts.ndim == 1:
            pts = pts.reshape(-1, 1)
        w = np.array(w, dtype=np.float64).ravel()
        if len(w) != pts.shape[0]:
            raise ValueError("number of weights and design points must be equal")
        if np.any(w < 0) or not math.isclose(w.sum(), 1.0, rel_tol=math.sqrt(np.finfo(float).eps)):
            raise ValueError("weights must be non-negative and sum to one")
        self.points = np.ascontiguousarray(pts)
        self.weights = w

    def __eq__(self, other):  # src/designmeasure.jl: ==
        return (isinstance(other, DesignMeasure) and np.array_equal(self.weights, other.weights)
                and np.array_equal(self.points, other.points))

    def __repr__(self):
        rows = ",\n ".join("%s => %r" % (list(p), float(w)) for p, w in zip(self.points, self.weights))
        return "DesignMeasure(\n %s,\n)" % rows

    def copy(self):
        return DesignMeasure(self.points.copy(), self.weights.copy())


def points(d):
    return [p for p in d.points]


def weights(d):
    return d.weights


def numpoints(d):
    return d.points.shape[0]


def one_point_design(designpoint):
    return DesignMeasure([designpoint], [1.0])


def uniform_design(designpoints):
    K = len(designpoints)
    return DesignMeasure(designpoints, [1.0 / K] * K)


def equidistant_design(dr, K):  # src/designmeasure.jl:162-169
    if dr.N != 1:
        raise TypeError("equidistant_design needs a 1-dimensional DesignInterval")
    if K <= 1:
        raise ValueError("equidistant design needs at least K = 2, got %d" % K)
    return uniform_design([[x] for x in np.linspace(dr.lowerbound[0], dr.upperbound[0], K)])


def random_design(dr, K, rng=None):  # src/designmeasure.jl:180-188
    rng = np.random.default_rng() if rng is None else rng
    lb, ub = np.array(dr.lowerbound), np.array(dr.upperbound)
    pts = lb + (ub - lb) * rng.random((K, dr.N))
    v = -np.log(rng.random(K))
    return DesignMeasure(pts, v / v.sum())


def sort_points(d, rev=False):  # src/designmeasure.jl: sort_points (stable lexicographic)
    idx = np.arange(numpoints(d))
    for i in range(d.points.shape[1] - 1, -1, -1):
        key = d.points[idx, i]
        order = np.argsort(-key if rev else key, kind="stable")
        idx = idx[order]
    return DesignMeasure(d.points[idx], d.weights[idx])


def sort_weights(d, rev=False):
    order = np.argsort(-d.weights if rev else d.weights, kind="stable")
    return DesignMeasure(d.points[order], d.weights[order])


def mixture(alpha, d1, d2):  # src/designmeasure.jl:312-322
    if d1.points.shape[1] != d2.points.shape[1]:
        raise ValueError("design points must have identical lengths")
    if alpha < 0 or alpha > 1:
        raise ValueError("mixture weight must be between 0 and 1")
    return DesignMeasure(np.vstack([d1.points, d2.points]),
                         np.concatenate([alpha * d1.weights, (1 - alpha) * d2.weights]))


def apportion(w, n):  # src/designmeasure.jl:335-351 (Pukelsheim 2006, p. 309)
    w = np.asarray(w.weights if isinstance(w, DesignMeasure) else w, dtype=float)
    a = np.ceil((n - 0.5 * len(w)) * w).astype(np.int64)
    while a.sum() != n:
        if a.sum() < n:
            a[np.argmin(a / w)] += 1
        else:
            a[np.argmax((a - 1) / w)] -= 1
    return a


def simplify_drop(d, maxweight):  # src/designmeasure.jl:372-381
    if numpoints(d) == 1:
        return d.copy()
    keep = d.weights > maxweight
    w = d.weights[keep]
    return DesignMeasure(d.points[keep], w / w.sum())


def simplify_unique(d, dr, m, cp, **uargs):  # src/designmeasure.jl:412-424 (generic catch-all)
    if uargs:
        import warnings
        warnings.warn("unused keyword arguments given to generic `simplify_unique` method: %r" % (uargs,))
    return d.copy()


def simplify_merge(d, dr, maxdist):  # src/designmeasure.jl:443-469
    if numpoints(d) == 1:
        return d.copy()
    width = np.array(dr.lowerbound) - np.array(dr.upperbound)
    dps = [p.copy() for p in d.points]
    ws = list(d.weights)
    cur = 0.0
    while cur <= maxdist:
        n = len(dps)
        if n < 2:
            break
        dist = np.full((n, n), np.inf)
        for i in range(n):
            for j in range(n):
                if i != j:
                    dist[i, j] = np.linalg.norm((dps[i] - dps[j]) / width)
        # Julia's findmin scans column-major: smallest column index first, then row
        flat = np.argmin(dist.T.ravel())
        j, i = divmod(flat, n)
        cur = dist[i, j]
        if cur <= maxdist:
            w_new = ws[i] + ws[j]
            dp_new = (ws[i] * dps[i] + ws[j] * dps[j]) / w_new
            keep = [k for k in range(n) if k != i and k != j]
            dps = [dp_new] + [dps[k] for k in keep]
            ws = [w_new] + [ws[k] for k in keep]
    return DesignMeasure(dps, ws)
