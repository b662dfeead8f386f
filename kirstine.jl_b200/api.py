"""Host-side mirror of Kirstine.jl's user-facing API for the accelerated path.

Same names, argument meaning and error behaviour as the reference (exported list:
src/Kirstine.jl:25-96), so parity tests read like the reference's own tests.  Everything that
evaluates an objective or a Gateaux derivative goes through the C ABI (include/kirstine_b200.h)
to the CUDA engine; there is no CPU fallback.  Differences forced by the host language:
indices are 0-based, keyword structs become Python classes, `DeltaMethod` takes a Python callable
that is evaluated on the host once per draw (exactly what trafo_constants does,
src/transformation.jl:71-92), and only *registered* model families (csrc/kst_models.cuh) can be
used -- arbitrary user `jacobianmatrix!` code cannot run on the GPU.
"""
import math
import time
import warnings
from collections import namedtuple

import numpy as np

from . import _ffi
from .designmeasure import (DesignInterval, DesignMeasure, mixture, numpoints, one_point_design, simplify_drop,
                            simplify_merge, simplify_unique, sort_points)

# ------------------------------------------------------------------------------------------------
# registered model families, their parameters and covariate parameterizations
# ------------------------------------------------------------------------------------------------
EmaxPar = namedtuple("EmaxPar", "e0 emax ec50")                    # test/example-emax.jl:12-16
SigEmaxParameter = namedtuple("SigEmaxParameter", "e0 emax ed50 h")  # docs/src/tutorial.md:116
TPCParameter = namedtuple("TPCParameter", "a e s")                  # test/example-compartment.jl:12
VUParameter = namedtuple("VUParameter", "a e s")                    # test/example-vector.jl:30
DTRParameter = namedtuple("DTRParameter", "a e e0 emax ec50")       # test/example-dtr.jl:22
SLRParameter = namedtuple("SLRParameter", "a b c")                  # docs/src/extend-model.md:342


class LLRParameter:
    """docs/src/extend-model.md:137-142: beta_0 plus a vector beta_rest; dimension = length(beta_rest) + 1."""

    def __init__(self, beta_0, beta_rest):
        self.beta_0 = float(beta_0)
        self.beta_rest = np.asarray(beta_rest, dtype=np.float64).ravel()

    def _flat(self):
        return np.concatenate([[self.beta_0], self.beta_rest])

    @classmethod
    def _from_flat(cls, row):
        return cls(row[0], row[1:])

    def __len__(self):
        return 1 + len(self.beta_rest)

    def __repr__(self):
        return "LLRParameter(%r, %r)" % (self.beta_0, list(self.beta_rest))


def dimension(p):
    return len(p)


class Model:
    model_id = None
    parameter_type = None
    m = 1

    def _params(self):
        raise NotImplementedError


class EmaxModel(Model):
    """test/example-emax.jl:4-6 -- note the field is sigma_squared."""
    model_id, parameter_type = _ffi.MODEL_EMAX, EmaxPar

    def __init__(self, sigma_squared):
        self.sigma_squared = float(sigma_squared)

    def _params(self):
        return [self.sigma_squared]


class SigEmaxModel(Model):
    """@simple_model SigEmax dose (docs/src/tutorial.md:107; src/user.jl:291-333)."""
    model_id, parameter_type = _ffi.MODEL_SIGEMAX, SigEmaxParameter

    def __init__(self, sigma):
        self.sigma = float(sigma)

    def _params(self):
        return [self.sigma]


class TPCModel(Model):
    """@simple_model TPC time (test/example-compartment.jl:10)."""
    model_id, parameter_type = _ffi.MODEL_TPC, TPCParameter

    def __init__(self, sigma):
        self.sigma = float(sigma)

    def _params(self):
        return [self.sigma]


class VUMod(Model):
    """test/example-vector.jl:9-12: vector units of length m, variance sigma_squared (1 + time)."""
    model_id, parameter_type = _ffi.MODEL_VU, VUParameter

    def __init__(self, sigma_squared, m):
        self.sigma_squared, self.m = float(sigma_squared), int(m)

    def _params(self):
        return [self.sigma_squared]


class DTRMod(Model):
    """test/example-dtr.jl:4-7, docs/src/dtr.md:87-90: dose-time-response, m measurements per unit."""
    model_id, parameter_type = _ffi.MODEL_DTR, DTRParameter

    def __init__(self, sigma, m):
        self.sigma, self.m = float(sigma), int(m)

    def _params(self):
        return [self.sigma]


class LogisticRegression(Model):
    """The user-defined Model supertype of docs/src/extend-model.md:59-106: unit length 1, no unit covariance,
    M = sum_k w_k p_k (1 - p_k) J_k' J_k with p = expit(nonlinear predictor), J = predictor Jacobian."""

    def _params(self):
        return [0.0]   # the C ABI wants at least one model parameter; logistic models ignore it


class LinLogReg(LogisticRegression):
    """docs/src/extend-model.md:123-179: linear predictor on a dim_covariate-dimensional covariate (1..3)."""
    model_id, parameter_type = _ffi.MODEL_LINLOGREG, LLRParameter

    def __init__(self, dim_covariate):
        self.dim_covariate = int(dim_covariate)


class SinLogReg(LogisticRegression):
    """docs/src/extend-model.md:330-355: predictor a sin(2 pi b (t - c))."""
    model_id, parameter_type = _ffi.MODEL_SINLOGREG, SLRParameter


def unit_length(m):
    return m.m


class CovariateParameterization:
    cp_id = _ffi.CP_COPY

    def _params(self):
        return []

    def simplify_unique(self, d, dr, m, **uargs):
        """Generic catch-all of the reference (src/designmeasure.jl:412-424)."""
        return simplify_unique(d, dr, m, self, **uargs)


class JustCopy(CovariateParameterization):
    """src/user.jl:107-109,185-216: covariate fields are copies of the design point."""

    def __init__(self, *field_names):
        self.field_names = field_names


class CopyDose(CovariateParameterization):  # test/example-emax.jl:18
    pass


class CopyVector(CovariateParameterization):  # docs/src/extend-model.md:145-155: covariate x .= design point
    pass


class CopyTime(CovariateParameterization):  # test/example-compartment.jl:16
    pass


class EquiTime(CovariateParameterization):  # test/example-vector.jl:43
    cp_id = _ffi.CP_EQUITIME


class CopyBoth(CovariateParameterization):
    """docs/src/dtr.md:197-203: dose = z[0], time = z[1]; with its simplify_unique (:229-247)."""
    cp_id = _ffi.CP_DTR_COPYBOTH

    def simplify_unique(self, d, dr, m, minposdose=0, minpostime=0):
        res = d.copy()
        if tuple(dr.lowerbound) == (0.0, 0.0):
            drop = (res.points[:, 0] < minposdose) | (res.points[:, 1] < minpostime)
            res.points[drop] = 0.0
        return res


class FixedDose(CovariateParameterization):  # docs/src/dtr.md:331-339
    cp_id = _ffi.CP_DTR_FIXEDDOSE

    def __init__(self, dose):
        self.dose = float(dose)

    def _params(self):
        return [self.dose]


class FixedTimes(CovariateParameterization):  # docs/src/dtr.md:401-409
    cp_id = _ffi.CP_DTR_FIXEDTIMES

    def __init__(self, time):
        self.time = [float(t) for t in time]

    def _params(self):
        return self.time


class EquidistantTimes(CovariateParameterization):
    """docs/src/dtr.md:484-495 with its simplify_unique (:498-523)."""
    cp_id = _ffi.CP_DTR_EQUIDISTANT

    def simplify_unique(self, d, dr, m, minposdose=0, minposdelta=0):
        res = d.copy()
        if dr.lowerbound[0] == 0:
            drop = res.points[:, 0] <= minposdose
            if dr.lowerbound[1] == 0:
                drop |= res.points[:, 1] <= minposdelta
            res.points[drop, 0] = 0.0
            res.points[drop, 1] = dr.upperbound[1]
        return res


class LogEquidistantTimes(EquidistantTimes):  # docs/src/dtr.md:663-681 (same canonical representative)
    cp_id = _ffi.CP_DTR_LOGEQUIDISTANT

    def __init__(self, b):
        self.b = float(b)

    def _params(self):
        return [self.b]


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
    sample split over several processes.
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
                h.comm_init(*self.comm)
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
        elif isinstance(m, VUMod):
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


def _solve_with_exchange(dp, strategy, trace_state):  # src/solve-exchange.jl:116-179
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
