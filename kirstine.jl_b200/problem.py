"""Criteria, transformations, normal approximations, prior samples, the design problem and its public evaluators
(objective, gateauxderivative, efficiency, ...); part of the API mirror described in the package docstring (__init__.py)."""
import math
import warnings

import numpy as np

from . import _ffi
from .designmeasure import (DesignMeasure, numpoints, one_point_design, simplify_drop, simplify_merge)
from .models import (DTRMod, EmaxModel, EquidistantTimes, FixedDose, FixedTimes, LinLogReg, LogEquidistantTimes, Model,
                     SigEmaxModel, SinLogReg, TPCModel, VUMod, VUModAR1)

# ------------------------------------------------------------------------------------------------
# criteria, transformations, normal approximation, prior knowledge
# ------------------------------------------------------------------------------------------------
class DCriterion:
    crit_id = _ffi.CRIT_D


class ACriterion:
    crit_id = _ffi.CRIT_A


class DexpCriterion:
    """det(M_T)^(1/t): the alternative D-criterion the reference's docs build as a user extension
    (docs/src/extend-criterion.md:100-183).  Not linear in log det, so Bayesian optima differ from DCriterion."""
    crit_id = _ffi.CRIT_DEXP


class Identity:
    pass


class DeltaMethod:
    """DeltaMethod(jacobian_matrix): parameter -> (t, r) Jacobian of the transformation."""

    def __init__(self, jacobian_matrix):
        self.jacobian_matrix = jacobian_matrix


class FisherMatrix:
    pass


class RegularizedFisherMatrix:
    """The user-extension normal approximation of docs/src/extend-approximation.md:69-79: informationmatrix!
    adds the constant matrix R to every average Fisher matrix."""

    def __init__(self, R):
        self.R = np.asarray(R, dtype=np.float64)


class PriorSample:
    """PriorSample(p[, weights]) (src/priorknowledge.jl:11-38)."""

    def __init__(self, parameters, weights=None):
        if len(parameters) == 0:
            raise ValueError("need at least one parameter value")
        types = {type(p) for p in parameters} if not isinstance(parameters, np.ndarray) else {np.ndarray}
        if len(types) > 1:
            raise ValueError("parameters must all have the same concrete type")
        p0 = parameters[0]
        self.ptype = type(p0) if hasattr(p0, "_fields") or hasattr(p0, "_flat") else None
        if hasattr(p0, "_flat"):
            parameters = [p._flat() for p in parameters]
        self.theta = np.ascontiguousarray(np.asarray(parameters, dtype=np.float64).reshape(len(parameters), -1))
        n = len(self.theta)
        w = np.full(n, 1.0 / n) if weights is None else np.asarray(weights, dtype=np.float64).ravel()
        if len(w) != n:
            raise ValueError("number of weights and parameter values must be equal")
        if np.any(w < 0) or not math.isclose(w.sum(), 1.0, rel_tol=math.sqrt(np.finfo(float).eps)):
            raise ValueError("weights must be non-negative and sum to one")
        self.weight = w

    def parameter(self, i):
        row = self.theta[i]
        if self.ptype is not None and hasattr(self.ptype, "_from_flat"):
            return self.ptype._from_flat(row)
        return self.ptype(*row) if self.ptype else row


def parameters(pk):
    return [pk.parameter(i) for i in range(len(pk.theta))]


def parameter_dimension(pk):
    return pk.theta.shape[1]


def trafo_constants(trafo, pk):
    """src/transformation.jl:71-92: None for Identity, (N, t, r) array for DeltaMethod."""
    if isinstance(trafo, Identity):
        return None
    jm = [np.atleast_2d(np.asarray(trafo.jacobian_matrix(pk.parameter(i)), dtype=np.float64))
          for i in range(len(pk.theta))]
    r = parameter_dimension(pk)
    if any(j.shape != jm[0].shape for j in jm):
        raise ValueError("trafo Jacobians must be identical in size")
    if jm[0].shape[1] != r:
        raise ValueError("trafo Jacobian must have %d columns, got %d" % (r, jm[0].shape[1]))
    if jm[0].shape[0] > jm[0].shape[1]:
        warnings.warn("trafo Jacobians have more rows than cols, infomatrices will be singular")
    elif any(np.linalg.matrix_rank(j) != jm[0].shape[0] for j in jm[: min(len(jm), 1000)]):
        warnings.warn("some trafo Jacobians are not full rank, infomatrices will be singular")
    return np.stack(jm)


def codomain_dimension(trafo, pk):
    if isinstance(trafo, Identity):
        return parameter_dimension(pk)
    return np.atleast_2d(np.asarray(trafo.jacobian_matrix(pk.parameter(0)))).shape[0]


# ------------------------------------------------------------------------------------------------
# design problem
# ------------------------------------------------------------------------------------------------
class DesignProblem:
    """DesignProblem(; criterion, region, model, covariate_parameterization, prior_knowledge,
    transformation = Identity(), normal_approximation = FisherMatrix()) (src/designproblem.jl:22-79).

    `device` / `comm` are engine options with no counterpart in the reference: the GPU to use and
    an optional (nranks, rank, unique_id) triple when the prior sample passed in is one shard of a
    sample split over several processes.  unique_id is the 128-byte NCCL id rank 0 made and every rank
    received, or a callable that does that exchange and returns a fresh id on every call: an id serves
    exactly one communicator, and efficiency / shannon_information on a non-D problem build a second
    handle (the same problem under DCriterion) that needs its own.
    """

    def __init__(self, *, criterion, region, model, covariate_parameterization, prior_knowledge,
                 transformation=None, normal_approximation=None, device=0, comm=None):
        self.dc, self.dr, self.m, self.cp, self.pk = criterion, region, model, covariate_parameterization, prior_knowledge
        self.trafo = Identity() if transformation is None else transformation
        self.na = FisherMatrix() if normal_approximation is None else normal_approximation
        self.device, self.comm = device, comm
        if not isinstance(model, Model) or model.model_id is None:
            raise TypeError("model type %s is not registered with the B200 engine; keep using stock Kirstine "
                            "for arbitrary user models" % type(model).__name__)
        if not isinstance(self.na, (FisherMatrix, RegularizedFisherMatrix)):
            raise TypeError("only the FisherMatrix and RegularizedFisherMatrix normal approximations are registered "
                            "with the B200 engine")
        self._handle_cache = None

    def handle(self):
        if self._handle_cache is None:
            dt = trafo_constants(self.trafo, self.pk)
            h = _ffi.ProblemHandle(self.m.model_id, self.m._params(), self.pk.theta, self.pk.weight, self.dc.crit_id,
                                   dt, self.dr.lowerbound, self.dr.upperbound, m=self.m.m, cp_id=self.cp.cp_id,
                                   device=self.device, cp_params=self.cp._params())
            if isinstance(self.na, RegularizedFisherMatrix):
                h.set_regularization(self.na.R)
            if self.comm is not None:
                nranks, rank, uid = self.comm
                h.comm_init(nranks, rank, uid() if callable(uid) else uid)
            self._handle_cache = h
        return self._handle_cache


def criterion(dp):
    return dp.dc


def region(dp):
    return dp.dr


def model(dp):
    return dp.m


def covariate_parameterization(dp):
    return dp.cp


def prior_knowledge(dp):
    return dp.pk


def transformation(dp):
    return dp.trafo


def normal_approximation(dp):
    return dp.na


def objective(d, dp):
    """objective(d, dp) (src/designproblem.jl:233-237)."""
    return float(dp.handle().objective_batch(d.points[None], d.weights[None])[0])


def gateauxderivative(at, directions, dp):
    """gateauxderivative(at, directions, dp) (src/designproblem.jl:273-297)."""
    directions = list(directions)
    if any(numpoints(d) != 1 for d in directions):
        raise RuntimeError("derivative only implemented for one-point directions")
    if math.isinf(objective(at, dp)):
        return np.full(len(directions), np.nan)
    h = dp.handle()
    if h.gateaux_prepare(at.points, at.weights):
        raise ArithmeticError("SingularException")
    return h.gateaux_eval(np.stack([d.points[0] for d in directions]))


def _as_doptimality_problem(dp):  # src/designproblem.jl:320-331
    if isinstance(dp.dc, DCriterion):
        return dp
    if getattr(dp, "_dopt", None) is None:
        if dp.comm is not None and not callable(dp.comm[2]):
            raise ValueError("the D-optimality twin of a multi-rank problem needs a communicator of its own: pass "
                             "comm=(nranks, rank, make_unique_id) with a callable that returns a fresh broadcast id")
        dp._dopt = DesignProblem(criterion=DCriterion(), region=dp.dr, model=dp.m, covariate_parameterization=dp.cp,
                                 prior_knowledge=dp.pk, transformation=dp.trafo, normal_approximation=dp.na,
                                 device=dp.device, comm=dp.comm)
    return dp._dopt


def efficiency(d1, d2, dp1, dp2=None):
    """Relative D-efficiency (src/designproblem.jl:315-350)."""
    dp2 = dp1 if dp2 is None else dp2
    t = codomain_dimension(dp1.trafo, dp1.pk)
    if t != codomain_dimension(dp2.trafo, dp2.pk):
        raise ValueError("dimensions of transformed parameters must match")  # DimensionMismatch
    with np.errstate(invalid="ignore", over="ignore"):
        return float(np.exp((np.float64(objective(d1, _as_doptimality_problem(dp1))) -
                             np.float64(objective(d2, _as_doptimality_problem(dp2)))) / t))


def shannon_information(d, dp, n):
    """src/designproblem.jl:366-370."""
    t = codomain_dimension(dp.trafo, dp.pk)
    return (t / 2) * (math.log(n) - 1 + math.log(2 * math.pi)) + 0.5 * objective(d, _as_doptimality_problem(dp))


def informationmatrix(d, m, cp, p, na):
    """informationmatrix(d, m, cp, p, na) for a single parameter value (src/design-common.jl:96-112)."""
    p = np.asarray(p, dtype=np.float64).reshape(1, -1)
    lo = np.minimum(d.points.min(axis=0), 0.0) - 1.0
    hi = d.points.max(axis=0) + 1.0
    h = _ffi.ProblemHandle(m.model_id, m._params(), p, [1.0], _ffi.CRIT_D, None, lo, hi, m=m.m, cp_id=cp.cp_id,
                           cp_params=cp._params())
    try:
        if isinstance(na, RegularizedFisherMatrix):
            h.set_regularization(na.R)
        return h.informationmatrix(d.points, d.weights, 0)
    finally:
        h.close()


def implied_covariates(d, m, cp):
    """implied_covariates(d, m, cp) (src/design-common.jl:35-49): the covariates a design implies, one dict per
    design point -- the host-side statement of what the device functors' prep_row does (a debugging helper, like
    in the reference)."""
    out = []
    for z in d.points:
        if isinstance(m, (EmaxModel, SigEmaxModel)):
            out.append({"dose": float(z[0])})
        elif isinstance(m, TPCModel):
            out.append({"time": float(z[0])})
        elif isinstance(m, (VUMod, VUModAR1)):
            out.append({"time": [float(j * z[0]) for j in range(m.m)]})
        elif isinstance(m, DTRMod):
            if isinstance(cp, FixedDose):
                dose, time = cp.dose, [float(z[0])]
            elif isinstance(cp, FixedTimes):
                dose, time = float(z[0]), list(cp.time)
            elif isinstance(cp, LogEquidistantTimes):
                time = [0.0] + ([float(z[1])] if m.m > 1 else [])
                for _ in range(2, m.m):
                    time.append(cp.b * time[-1])
                dose = float(z[0])
            elif isinstance(cp, EquidistantTimes):
                dose, time = float(z[0]), [float(j * z[1]) for j in range(m.m)]
            else:
                dose, time = float(z[0]), [float(z[1])]
            out.append({"dose": dose, "time": time})
        elif isinstance(m, LinLogReg):
            if len(z) != m.dim_covariate:
                raise ValueError("dimension of covariate and design point don't match")
            out.append({"x": [float(v) for v in z]})
        elif isinstance(m, SinLogReg):
            out.append({"t": float(z[0])})
        else:
            raise TypeError("model type %s is not registered with the B200 engine" % type(m).__name__)
    return out


def apply_transformation(M, p, trafo):
    """apply_transformation(nim, p, trafo) (src/design-common.jl:113-142): the information matrix of the
    transformed parameter for ONE matrix and ONE parameter value, M_T = (DT M^-1 DT')^-1 (Identity: M itself).
    A debugging helper of the reference on a single r x r matrix; the engine evaluates the same expression for
    every (particle, draw) inside kst_criterion.  Only the upper triangle of M is read."""
    M = np.asarray(M, dtype=np.float64)
    M = np.triu(M) + np.triu(M, 1).T
    if isinstance(trafo, Identity):
        return M.copy()
    DT = np.atleast_2d(np.asarray(trafo.jacobian_matrix(p), dtype=np.float64))
    if DT.shape[1] != M.shape[0]:
        raise ValueError("trafo Jacobian must have %d columns, got %d" % (M.shape[0], DT.shape[1]))
    L = np.linalg.cholesky(M)                      # PosDefException in the reference when M is not PD
    B = np.linalg.solve(L.T, np.linalg.solve(L, DT.T))
    return np.linalg.inv(DT @ B)


def gateauxderivative_grid(d, dp, n_grid=None, csv_path=None):
    """The data plot_gateauxderivative draws (src/plot.jl:103-187), without a plotting backend: the Gateaux
    derivative of `d` on a grid over the bounding box of the design region.

    1-d regions: n_grid equidistant points (default 101) merged with the design's own points and sorted; returns
    (x_grid, gd).  2-d regions: an n_grid[0] x n_grid[1] product grid (default 51 x 51); returns (range_x, range_y,
    gd) with gd[i, j] at (range_x[i], range_y[j]).  Points outside the region would be NaN (every DesignInterval
    point of its bounding box is inside).  csv_path: also write the grid as CSV (columns = coordinates, gd)."""
    lb, ub = np.asarray(dp.dr.lowerbound), np.asarray(dp.dr.upperbound)
    dim = len(lb)
    if dim == 1:
        n = 101 if n_grid is None else n_grid
        if np.ndim(n) != 0:
            raise ValueError("n_grid must be a single integer")
        if n < 2:
            raise ValueError("n_grid must be >= 2")
        x = np.sort(np.concatenate([np.linspace(lb[0], ub[0], int(n)), d.points.ravel()]))
        gd = gateauxderivative(d, [one_point_design([v]) for v in x], dp)
        out = (x, gd)
        table = np.column_stack([x, gd])
    elif dim == 2:
        n = (51, 51) if n_grid is None else tuple(np.atleast_1d(n_grid).tolist())
        if len(n) != 2:
            raise ValueError("n_grid must be a 2-tuple")
        if min(n) < 2:
            raise ValueError("all elements of n_grid must be >= 2")
        rx, ry = np.linspace(lb[0], ub[0], n[0]), np.linspace(lb[1], ub[1], n[1])
        gx, gy = np.meshgrid(rx, ry, indexing="ij")
        gd = gateauxderivative(d, [one_point_design([a, b]) for a, b in zip(gx.ravel(), gy.ravel())], dp).reshape(n)
        out = (rx, ry, gd)
        table = np.column_stack([gx.ravel(), gy.ravel(), gd.ravel()])
    else:
        raise ValueError("grids are defined for 1- and 2-dimensional design regions")
    if csv_path is not None:
        np.savetxt(csv_path, table, delimiter=",", header=",".join(list(dp.dr.name) + ["gateauxderivative"]), comments="")
    return out


def simplify(d, dp, maxweight=0, maxdist=0, **uargs):
    """src/designproblem.jl:185-190."""
    d = dp.cp.simplify_unique(d, dp.dr, dp.m, **uargs)
    d = simplify_merge(d, dp.dr, maxdist)
    d = simplify_drop(d, maxweight)
    return d
