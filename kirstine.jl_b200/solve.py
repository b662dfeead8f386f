"""Optimizer (device-resident PSO), problem solving strategies and their results; part of the API mirror described
in the package docstring (__init__.py)."""
import math
import time
import warnings

import numpy as np

from . import _ffi
from .designmeasure import DesignMeasure, mixture, numpoints, one_point_design, simplify_merge, sort_points
from .models import CovariateParameterization
from .problem import objective, simplify

# ------------------------------------------------------------------------------------------------
# optimizer, strategies, results
# ------------------------------------------------------------------------------------------------
class Pso:
    """Pso(; iterations, swarmsize, c1 = 2.05, c2 = 2.05) (src/pso.jl:29-59)."""

    def __init__(self, *, iterations, swarmsize, c1=2.05, c2=2.05):
        if iterations <= 0:
            raise ValueError("number of iterations must be positive")
        if swarmsize <= 1:
            raise ValueError("swarmsize must be at least 2")
        if c1 < 0 or c2 < 0 or c1 + c2 <= 4:
            raise ValueError("c1 and c2 must be non-negative and, c1 + c2 > 4")
        self.iterations, self.swarmsize, self.c1, self.c2 = int(iterations), int(swarmsize), float(c1), float(c2)


class PsoState:
    """PsoState (src/pso.jl:63-76) as copied out of the device.  x, p (lists of DesignMeasure) and g are built from
    the raw arrays on first access: wrapping a whole swarm in Python objects costs more than a short PSO run."""

    def __init__(self, st):
        self.v_atoms, self.v_weights = st["v_pts"], st["v_w"]
        self.fx, self.fp, self.fg, self.n_eval = st["fx"], st["fp"], st["fg"], st["n_eval"]
        self.raw = st
        self._x = self._p = self._g = None

    @property
    def x(self):
        if self._x is None:
            self._x = [DesignMeasure(self.raw["x_pts"][i], _renorm(self.raw["x_w"][i])) for i in range(len(self.fx))]
        return self._x

    @property
    def p(self):
        if self._p is None:
            self._p = [DesignMeasure(self.raw["p_pts"][i], _renorm(self.raw["p_w"][i])) for i in range(len(self.fx))]
        return self._p

    @property
    def g(self):
        if self._g is None:
            self._g = DesignMeasure(self.raw["g_pts"], _renorm(self.raw["g_w"]))
        return self._g


def _renorm(w):
    # weights coming back from the swarm sum to one only up to rounding
    return np.asarray(w, dtype=np.float64)


class OptimizationResult:
    """OptimizationResult (src/optimize.jl:28-40)."""

    def __init__(self, maximizer, maximum, trace_x, trace_fx, trace_state, n_eval, seconds_elapsed):
        self.maximizer, self.maximum = maximizer, maximum
        self.trace_x, self.trace_fx, self.trace_state = trace_x, trace_fx, trace_state
        self.n_eval, self.seconds_elapsed = n_eval, seconds_elapsed

    def __repr__(self):
        return "OptimizationResult(maximum: %r, iterations: %d, evaluations: %d)" % (
            self.maximum, len(self.trace_fx), self.n_eval)


_seed_counter = [20231107]


def seed(s):
    """Stand-in for Random.seed!: seeds the engine's Philox stream for the following solves."""
    _seed_counter[0] = int(s)


def _next_seed():
    s = _seed_counter[0]
    _seed_counter[0] = (s * 6364136223846793005 + 1442695040888963407) % (1 << 64)
    return s


def _check_compatible(d, dr):  # src/designmeasure-abstractpoint.jl:60-71
    if d.points.shape[1] != dr.N:
        raise ValueError("design point length must match design region dimension")
    if np.any(d.points < np.array(dr.lowerbound)) or np.any(d.points > np.array(dr.upperbound)):
        raise ValueError("design point is outside design region")


def _design_constraints(d, dr, fixedweights, fixedpoints):  # src/designmeasure-abstractpoint.jl:35-58
    _check_compatible(d, dr)
    K = numpoints(d)
    fixedweights, fixedpoints = list(fixedweights), list(fixedpoints)
    if any(k < 0 or k >= K for k in fixedweights):
        raise ValueError("indices for fixed weights must be between 0 and %d" % (K - 1))
    if any(k < 0 or k >= K for k in fixedpoints):
        raise ValueError("indices for fixed points must be between 0 and %d" % (K - 1))
    fixw = np.array([k in fixedweights for k in range(K)], dtype=np.int32)
    fixp = np.array([k in fixedpoints for k in range(K)], dtype=np.int32)
    if fixw.sum() == K - 1:
        fixw[:] = 1  # "explicitly fixing implicitly fixed weight"
    return fixw, fixp


def _wrap_design(pts, w):
    dm = DesignMeasure.__new__(DesignMeasure)
    dm.points = np.ascontiguousarray(pts, dtype=np.float64)
    dm.weights = np.asarray(w, dtype=np.float64)
    return dm


def optimize(h, optimizer, prototypes, fixw, fixp, mode, trace_state=False):
    """optimize (src/optimize.jl:42-70) with the whole swarm resident on the device."""
    t0 = time.time()
    K = numpoints(prototypes[0])
    if not trace_state:
        # the whole run in one call: nothing but the traces, the status and the initial state's snapshot come back
        st, fx, px, pw, s0 = h.pso_optimize(optimizer.swarmsize, K, optimizer.iterations,
                                            np.stack([p.points for p in prototypes]),
                                            np.stack([p.weights for p in prototypes]), fixw, fixp, seed=_next_seed(),
                                            c1=optimizer.c1, c2=optimizer.c2, mode=mode)
        if st == 1:
            raise RuntimeError("infinite objective value at best initial point")
        if st == 2:
            raise RuntimeError("objective evaluates to NaN on all initial points")
        t_x = [_wrap_design(px[i], pw[i]) for i in range(optimizer.iterations)]
        fxm = float(fx[-1])
        if not math.isfinite(fxm):
            warnings.warn("maximum is not finite")
        return OptimizationResult(t_x[-1], fxm, t_x, [float(v) for v in fx], [PsoState(s0)],
                                  optimizer.swarmsize * optimizer.iterations, time.time() - t0)
    st = h.pso_init(optimizer.swarmsize, K, np.stack([p.points for p in prototypes]),
                    np.stack([p.weights for p in prototypes]), fixw, fixp, seed=_next_seed(), c1=optimizer.c1,
                    c2=optimizer.c2, mode=mode)
    if st == 1:
        raise RuntimeError("infinite objective value at best initial point")
    if st == 2:
        raise RuntimeError("objective evaluates to NaN on all initial points")
    s0 = h.pso_state()
    t_x = [_wrap_design(s0["g_pts"], s0["g_w"])]
    t_fx = [s0["fg"]]
    t_state = [PsoState(s0)]
    n = optimizer.iterations - 1
    if trace_state:
        for _ in range(n):
            fx, px, pw = h.pso_run(1)
            t_x.append(_wrap_design(px[0], pw[0]))
            t_fx.append(float(fx[0]))
            t_state.append(PsoState(h.pso_state()))
    elif n > 0:
        fx, px, pw = h.pso_run(n)
        t_x += [_wrap_design(px[i], pw[i]) for i in range(n)]
        t_fx += [float(v) for v in fx]
    final = h.pso_state()
    x = _wrap_design(final["g_pts"], final["g_w"])
    fxm = final["fg"]
    if not math.isfinite(fxm):
        warnings.warn("maximum is not finite")
    return OptimizationResult(x, fxm, t_x, t_fx, t_state, final["n_eval"], time.time() - t0)


class DirectMaximization:
    """DirectMaximization(; optimizer, prototype, fixedweights = [], fixedpoints = [])
    (src/solve-directmaximization.jl:11-48); indices are 0-based here."""

    def __init__(self, *, optimizer, prototype, fixedweights=(), fixedpoints=()):
        self.optimizer, self.prototype = optimizer, prototype
        self.fixedweights, self.fixedpoints = list(fixedweights), list(fixedpoints)


class DirectMaximizationResult:
    def __init__(self, or_):
        self.or_ = or_


class Exchange:
    """Exchange(; optimizer_direction, optimizer_weight, steps, candidate, simplify_args = {})
    (src/solve-exchange.jl:15-69)."""

    def __init__(self, *, optimizer_direction, optimizer_weight, steps, candidate, simplify_args=None):
        if steps < 1:
            raise ValueError("steps must be >= 1")
        self.optimizer_direction, self.optimizer_weight = optimizer_direction, optimizer_weight
        self.steps, self.candidate = int(steps), candidate
        self.simplify_args = dict(simplify_args or {})


class ExchangeResult:
    def __init__(self, ord_, orw):
        self.ord, self.orw = ord_, orw


def solution(res):
    if isinstance(res, DirectMaximizationResult):
        return res.or_.maximizer
    return res.orw[-1].maximizer


def optimization_result(dmr):
    return dmr.or_


def optimization_results_direction(er):
    return er.ord


def optimization_results_weight(er):
    return er.orw


def _solve_with_direct(dp, strategy, trace_state):  # src/solve-directmaximization.jl:80-97
    if not isinstance(strategy.optimizer, Pso):
        raise TypeError("the B200 engine runs Pso only")
    fixw, fixp = _design_constraints(strategy.prototype, dp.dr, strategy.fixedweights, strategy.fixedpoints)
    or_ = optimize(dp.handle(), strategy.optimizer, [strategy.prototype], fixw, fixp, _ffi.PSO_MODE_OBJECTIVE,
                   trace_state)
    return DirectMaximizationResult(or_)


def _exchange_unique_args(dp, simplify_args):
    """The keyword arguments a registered covariate parameterization's simplify_unique takes, as the two doubles of
    kst_exchange_options; None when the arguments cannot be expressed that way (the caller then steps in Python)."""
    from .models import CopyBoth, EquidistantTimes
    extra = {k: v for k, v in simplify_args.items() if k not in ("maxweight", "maxdist")}
    if isinstance(dp.cp, CopyBoth):
        names = ("minposdose", "minpostime")
    elif isinstance(dp.cp, EquidistantTimes):
        names = ("minposdose", "minposdelta")
    else:
        return (0.0, 0.0) if not extra and type(dp.cp).simplify_unique is CovariateParameterization.simplify_unique else None
    if any(k not in names for k in extra):
        return None
    return float(extra.get(names[0], 0)), float(extra.get(names[1], 0))


def _solve_with_exchange_device(dp, strategy, ua_ub):
    """The whole Exchange loop behind kst_exchange_run (one C-ABI call, both optimizers on the device)."""
    candidate = strategy.candidate
    _check_compatible(candidate, dp.dr)
    od, ow = strategy.optimizer_direction, strategy.optimizer_weight
    seed0 = _seed_counter[0]
    for _ in range(2 * strategy.steps):     # the call consumes the same seeds as 2 * steps optimize() calls would
        _next_seed()
    t0 = time.time()
    try:
        r = dp.handle().exchange_run(candidate.points, candidate.weights, strategy.steps,
                                     (od.swarmsize, od.iterations, od.c1, od.c2), (ow.swarmsize, ow.iterations, ow.c1, ow.c2),
                                     maxweight=strategy.simplify_args.get("maxweight", 0),
                                     maxdist=strategy.simplify_args.get("maxdist", 0), unique_a=ua_ub[0], unique_b=ua_ub[1],
                                     seed=seed0)
    except _ffi.EngineError as exc:
        if exc.code == -7:
            raise ArithmeticError("SingularException") from exc
        if exc.code == -8:
            raise RuntimeError(str(exc).split(": ", 1)[-1]) from exc
        raise
    secs = (time.time() - t0) / (2 * strategy.steps)
    ors_d, ors_w = [], []
    for s_ in range(strategy.steps):
        tx = [_wrap_design(r["dir_x"][s_, i].reshape(1, -1), [1.0]) for i in range(od.iterations)]
        fx = [float(v) for v in r["dir_fx"][s_]]
        ors_d.append(OptimizationResult(tx[-1], fx[-1], tx, fx, [PsoState(r["dir_state0"][s_])],
                                        od.swarmsize * od.iterations, secs))
        tx = [_wrap_design(r["w_points"][s_][i], r["w_weights"][s_][i]) for i in range(ow.iterations)]
        fx = [float(v) for v in r["w_fx"][s_]]
        if not math.isfinite(fx[-1]):
            warnings.warn("maximum is not finite")
        ors_w.append(OptimizationResult(tx[-1], fx[-1], tx, fx, [PsoState(r["w_state0"][s_])],
                                        ow.swarmsize * ow.iterations, secs))
    return ExchangeResult(ors_d, ors_w)


def _solve_with_exchange(dp, strategy, trace_state):  # src/solve-exchange.jl:116-179
    if not trace_state and isinstance(strategy.optimizer_direction, Pso) and isinstance(strategy.optimizer_weight, Pso):
        ua_ub = _exchange_unique_args(dp, strategy.simplify_args)
        if ua_ub is not None:
            return _solve_with_exchange_device(dp, strategy, ua_ub)
    # trace_state = True (every iteration's swarm is wanted on the host) or a simplify_unique the library does not
    # know: the same loop, one C-ABI call per optimizer run
    candidate = strategy.candidate
    _check_compatible(candidate, dp.dr)
    h = dp.handle()
    ors_d, ors_w = [], []
    for _ in range(strategy.steps):
        candidate = simplify(candidate, dp, **strategy.simplify_args)
        prototypes = [one_point_design(p) for p in candidate.points]
        if h.gateaux_prepare(candidate.points, candidate.weights):
            raise ArithmeticError("SingularException")
        or_gd = optimize(h, strategy.optimizer_direction, prototypes, [0], [0], _ffi.PSO_MODE_GATEAUX, trace_state)
        direction = or_gd.maximizer
        if any(np.array_equal(direction.points[0], p) for p in candidate.points):
            candidate = mixture(0, direction, candidate)
            candidate = simplify_merge(candidate, dp.dr, 0)
        else:
            candidate = mixture(1 / (1 + numpoints(candidate)), direction, candidate)
        wstr = DirectMaximization(optimizer=strategy.optimizer_weight, prototype=candidate,
                                  fixedpoints=range(numpoints(candidate)))
        _, rw = solve(dp, wstr, trace_state=trace_state)
        candidate = solution(rw)
        ors_d.append(or_gd)
        ors_w.append(optimization_result(rw))
    return ExchangeResult(ors_d, ors_w)


def solve(dp, strategy, trace_state=False, **sargs):
    """solve(dp, strategy; trace_state = false, sargs...) -> (design, result)
    (src/designproblem.jl:159-175)."""
    if isinstance(strategy, DirectMaximization):
        res = _solve_with_direct(dp, strategy, trace_state)
    elif isinstance(strategy, Exchange):
        res = _solve_with_exchange(dp, strategy, trace_state)
    else:
        raise TypeError("unknown problem solving strategy %r" % (strategy,))
    sol = solution(res)
    sol = DesignMeasure(sol.points, sol.weights / sol.weights.sum())
    dopt = sort_points(simplify(sol, dp, **sargs))
    o_before = objective(sol, dp)
    o_after = objective(dopt, dp)
    rel_diff = (o_after - o_before) / abs(o_before)
    if rel_diff < -0.01:
        warnings.warn("simplification may have been too eager (o_before=%r, o_after=%r)" % (o_before, o_after))
    return dopt, res
