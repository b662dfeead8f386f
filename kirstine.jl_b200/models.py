"""Registered model families, their parameter types and covariate parameterizations (host side of
csrc/kst_models.cuh); part of the API mirror described in the package docstring (__init__.py)."""
from collections import namedtuple

import numpy as np

from . import _ffi
from .designmeasure import simplify_unique

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


class VUModAR1(Model):
    """VUMod (test/example-vector.jl:9-12) with AR(1)-correlated errors within a unit:
    Sigma_ij = sigma_squared rho^|i-j| sqrt((1 + t_i)(1 + t_j)).  The registered example of a NonlinearRegression whose
    update_model_vcov! fills a dense matrix (the reference factors any Sigma(c) with potrf! and whitens with trsm!,
    src/model-nonlinear-regression.jl:39-49,79-83); rho = 0 is VUMod."""
    model_id, parameter_type = _ffi.MODEL_VUAR, VUParameter

    def __init__(self, sigma_squared, rho, m):
        self.sigma_squared, self.rho, self.m = float(sigma_squared), float(rho), int(m)

    def _params(self):
        return [self.sigma_squared, self.rho]


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
