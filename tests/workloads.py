"""Seeded synthetic inputs for the BASELINE.json configs, shared by tests and bench.py."""
import numpy as np


def emax_prior(n, seed=1):
    rng = np.random.default_rng(seed)
    return np.column_stack([rng.normal(1, 0.5, n), rng.normal(10, 2, n), np.exp(rng.normal(np.log(5), 0.3, n))])


def sigemax_prior(n, seed=31415):  # docs/src/tutorial.md:190-196
    rng = np.random.default_rng(seed)
    return np.array([1, 2, 0.4, 5]) + np.array([0.5, 0.5, 0.05, 0.5]) * rng.standard_normal((n, 4))


def tpc_prior(n, seed=3, se_factor=2):  # test/example-compartment.jl:74-82 (draw_from_prior)
    rng = np.random.default_rng(seed)
    a = 4.298 + se_factor * 0.5 * (2 * rng.random(n) - 1)
    e = 0.05884 + se_factor * 0.005 * (2 * rng.random(n) - 1)
    return np.column_stack([a, e, np.full(n, 21.8)])


def ded90_jacobians(theta, rho=90.0):  # docs/src/autodiff.md:257-266 (DED_manual)
    a = (100 / rho - 1) ** (1 / theta[:, 3])
    dt = np.zeros((len(theta), 1, 4))
    dt[:, 0, 2] = 1 / a
    dt[:, 0, 3] = theta[:, 2] * np.log(100 / rho - 1) / (theta[:, 3] ** 2 * a)
    return dt


def random_swarm(S, K, lb, ub, seed=0, prototype=None):
    """Particle 0 = equidistant prototype, others uniform points + Exp(1)-normalised weights
    (the reference's own initialiser, src/designmeasure-abstractpoint.jl:73-107)."""
    rng = np.random.default_rng(seed)
    lb, ub = np.atleast_1d(np.asarray(lb, float)), np.atleast_1d(np.asarray(ub, float))
    d = len(lb)
    pts = lb + (ub - lb) * rng.random((S, K, d))
    v = -np.log(1.0 - rng.random((S, K)))
    w = v / v.sum(axis=1, keepdims=True)
    if prototype is None:
        pts[0] = np.linspace(lb, ub, K).reshape(K, d)
        w[0] = 1.0 / K
    else:
        pts[0], w[0] = prototype
    return pts, w


def jittered_swarm(S, K, lb, ub, seed=0, log_lo=None):
    """Well-separated random designs for parity tests: point k is drawn inside the k-th of K equal
    cells (of [lb, ub], or of [log_lo, ub] on the log scale), weights are Dirichlet(4,...,4).
    Random i.i.d. points (random_swarm) are often numerically singular, where even the reference's
    own arithmetic is rounding noise."""
    rng = np.random.default_rng(seed)
    u = (np.arange(K) + 0.15 + 0.7 * rng.random((S, K))) / K
    if log_lo is None:
        pts = lb + (ub - lb) * u
    else:
        pts = np.exp(np.log(log_lo) + u * (np.log(ub) - np.log(log_lo)))
    w = rng.dirichlet(np.full(K, 4.0), size=S)
    return pts.reshape(S, K, 1), w


def dtr_prior(n, seed=4711):  # docs/src/dtr.md:133-147 (draw_from_prior)
    rng = np.random.default_rng(seed)
    return np.column_stack([0.5 + 0.10 * rng.standard_normal(n), 0.1 + 0.02 * rng.standard_normal(n),
                            1.0 + 0.25 * rng.standard_normal(n), 2.0 + 0.50 * rng.standard_normal(n),
                            np.exp(2.3 + 0.6 * rng.standard_normal(n))])


def jittered_swarm_nd(S, K, lb, ub, seed=0):
    """Well-separated random designs in a d-dimensional box: coordinate c of point k lies in the
    ((k * (c + 1)) mod K)-th of K equal cells, so no two points share a cell in any coordinate."""
    rng = np.random.default_rng(seed)
    lb, ub = np.asarray(lb, float), np.asarray(ub, float)
    d = len(lb)
    pts = np.empty((S, K, d))
    for c in range(d):
        cell = (np.arange(K) * (c + 1) + c) % K
        u = (cell + 0.15 + 0.7 * rng.random((S, K))) / K
        pts[:, :, c] = lb[c] + (ub[c] - lb[c]) * u
    w = rng.dirichlet(np.full(K, 4.0), size=S)
    return pts, w
