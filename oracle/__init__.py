"""ctypes front-end of the CPU parity oracle (oracle/kirstine_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Nothing under the product
package imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libkirstine_oracle.so")

EMAX, SIGEMAX, TPC, VU, DTR, LINLOGREG, SINLOGREG, VUAR = 1, 2, 3, 4, 5, 6, 7, 8
CP_COPY, CP_EQUITIME = 0, 1
CP_DTR_COPYBOTH, CP_DTR_FIXEDDOSE, CP_DTR_FIXEDTIMES, CP_DTR_EQUIDISTANT, CP_DTR_LOGEQUIDISTANT = 2, 3, 4, 5, 6
CRIT_D, CRIT_A, CRIT_DEXP = 0, 1, 2
TRAFO_IDENTITY, TRAFO_DELTA = 0, 1


def build(force=False):
    src = os.path.join(_HERE, "kirstine_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libkirstine_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.ko_create.restype = C.c_void_p
        L.ko_create.argtypes = [C.c_int, _dp, C.c_int, C.c_int, _dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_int, C.c_int, C.c_long, _dp, _dp, _dp]
        L.ko_destroy.argtypes = [C.c_void_p]
        L.ko_set_regularization.argtypes = [C.c_void_p, _dp]
        L.ko_objective.restype = C.c_double
        L.ko_objective.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
        L.ko_objective_range.restype = C.c_double
        L.ko_objective_range.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.c_long, C.c_long]
        L.ko_objective_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp, C.c_int]
        L.ko_informationmatrix.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.c_long, _dp]
        L.ko_gateaux_constants.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp, _dp]
        L.ko_gateaux_batch.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, _dp, C.c_int]
        L.ko_jacobian.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.ko_log_det.restype = C.c_double
        L.ko_log_det.argtypes = [C.c_int, _dp]
        L.ko_potrf_upper.argtypes = [C.c_int, _dp]
        L.ko_tr_prod_upper.restype = C.c_double
        L.ko_tr_prod_upper.argtypes = [C.c_int, _dp, _dp]
        L.ko_criterion_functional.restype = C.c_double
        L.ko_criterion_functional.argtypes = [C.c_int, _dp, C.c_int, C.c_int, _ip]
        L.ko_apply_transformation.argtypes = [C.c_int, C.c_int, C.c_int, _dp, C.c_int, _dp, _dp, _ip]
        L.ko_move_designpoint_how_far.restype = C.c_double
        L.ko_move_designpoint_how_far.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
        L.ko_move_handle_fixed.argtypes = [C.c_int, C.c_int, _dp, _dp, _ip, _ip]
        L.ko_move_how_far.restype = C.c_double
        L.ko_move_how_far.argtypes = [C.c_int, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
        L.ko_move_add_v.argtypes = [C.c_int, C.c_int, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _ip]
        L.ko_ap_add.restype = C.c_double
        L.ko_ap_add.argtypes = [C.c_int, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp, _ip, _ip]
        L.ko_philox4x32.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                    C.POINTER(C.c_uint32)]
        L.ko_uniform2_export.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _dp]
        L.ko_pso_create.restype = C.c_void_p
        L.ko_pso_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, _dp, _dp,
                                    _ip, _ip, C.c_uint64, C.c_int, _dp, _dp, C.c_int]
        L.ko_pso_destroy.argtypes = [C.c_void_p]
        L.ko_pso_init.argtypes = [C.c_void_p, _dp, _dp, C.c_int]
        L.ko_pso_tick.argtypes = [C.c_void_p]
        L.ko_pso_get_state.argtypes = [C.c_void_p] + [_dp] * 11 + [C.POINTER(C.c_long)]
        _lib = L
    return _lib


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def _i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


class Problem:
    """One design problem on the oracle side: model, criterion, transformation, prior sample.

    theta: (N, r) array, one parameter struct per row.  dt: (N, t, r) DeltaMethod
    Jacobians or None.  points passed to the methods are (K, d) or (S, K, d).
    """

    def __init__(self, model_id, model_params, theta, prior_weights=None, crit=CRIT_D,
                 dt=None, m=1, d=1, cp_id=CP_COPY, cp_params=()):
        theta = _f64(theta)
        self.N, self.r = theta.shape
        self.m, self.d = int(m), int(d)
        pw = _f64(np.full(self.N, 1.0 / self.N) if prior_weights is None else prior_weights)
        self.trafo = TRAFO_IDENTITY if dt is None else TRAFO_DELTA
        if dt is None:
            self.t = self.r
            dtc = None
        else:
            dt = _f64(dt)
            self.t = dt.shape[1]
            # per draw column-major t x r, like Julia's Matrix{Float64}
            dtc = _f64(np.transpose(dt, (0, 2, 1)).reshape(self.N, -1))
        mp = _f64(model_params)
        cpp = _f64(cp_params) if len(cp_params) else None
        self.crit = crit
        self.model_id = model_id
        self._h = lib().ko_create(model_id, _d(mp), len(mp), cp_id, _d(cpp), 0 if cpp is None else len(cpp), crit,
                                  self.trafo, self.r, self.t,
                                  self.m, self.d, self.N, _d(theta), _d(pw), _d(dtc))
        if not self._h:
            raise ValueError("oracle: dimensions out of range")
        self.theta, self.pw, self.dt = theta, pw, dt

    def set_regularization(self, R):
        """normal_approximation = RegularizedFisherMatrix(R) (docs/src/extend-approximation.md:69-79); None = FisherMatrix."""
        Rc = None if R is None else _f64(np.asarray(R, dtype=np.float64).T)   # column-major
        lib().ko_set_regularization(self._h, _d(Rc))
        return self

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.ko_destroy(self._h)
            self._h = None

    def _pw(self, points, weights):
        pts = _f64(points).reshape(-1, self.d) if np.ndim(points) <= 2 else _f64(points)
        return pts, _f64(weights)

    def objective(self, points, weights):
        pts = _f64(points).reshape(-1, self.d)
        w = _f64(weights)
        return lib().ko_objective(self._h, len(w), _d(pts), _d(w))

    def objective_range(self, points, weights, j0, j1):
        pts = _f64(points).reshape(-1, self.d)
        w = _f64(weights)
        return lib().ko_objective_range(self._h, len(w), _d(pts), _d(w), j0, j1)

    def objective_batch(self, points, weights, nthreads=1):
        w = _f64(weights)
        S, K = w.shape
        pts = _f64(points).reshape(S, K, self.d)
        fx = np.empty(S)
        lib().ko_objective_batch(self._h, S, K, _d(pts), _d(w), _d(fx), nthreads)
        return fx

    def informationmatrix(self, points, weights, draw=0):
        pts = _f64(points).reshape(-1, self.d)
        w = _f64(weights)
        out = np.empty((self.r, self.r))
        rc = lib().ko_informationmatrix(self._h, len(w), _d(pts), _d(w), draw, _d(out))
        assert rc == 0
        return out

    def gateaux_constants(self, points, weights):
        pts = _f64(points).reshape(-1, self.d)
        w = _f64(weights)
        A = np.empty((self.N, self.r, self.r))
        trB = np.empty(self.N)
        rc = lib().ko_gateaux_constants(self._h, len(w), _d(pts), _d(w), _d(A), _d(trB))
        if rc:
            raise ArithmeticError("SingularException")
        # stored column-major per draw; A_j is symmetric up to rounding, return row-major view
        return np.transpose(A, (0, 2, 1)).copy(), trB

    def gateauxderivative(self, at_points, at_weights, directions, nthreads=1):
        """Public gateauxderivative (src/designproblem.jl:273-297): NaN fill for singular `at`."""
        dirs = _f64(directions).reshape(-1, self.d)
        if np.isinf(self.objective(at_points, at_weights)):
            return np.full(len(dirs), np.nan)
        A, trB = self.gateaux_constants(at_points, at_weights)
        return self.gateaux_batch(dirs, A, trB, nthreads)

    def gateaux_batch(self, directions, A, trB, nthreads=1):
        dirs = _f64(directions).reshape(-1, self.d)
        Ac = _f64(np.transpose(A, (0, 2, 1)))
        trB = _f64(trB)
        out = np.empty(len(dirs))
        lib().ko_gateaux_batch(self._h, len(dirs), _d(dirs), _d(Ac), _d(trB), _d(out), nthreads)
        return out

    def jacobian(self, covariate, theta):
        c = _f64(covariate)
        th = _f64(theta)
        jm = np.empty((self.r, self.m))
        lib().ko_jacobian(self._h, _d(c), _d(th), _d(jm))
        return jm.T.copy()


def log_det(A):
    A = np.asfortranarray(_f64(A)).copy(order="F")
    return lib().ko_log_det(A.shape[0], A.ctypes.data_as(_dp))


def tr_prod_upper(A, B):
    A = np.asfortranarray(_f64(A))
    B = np.asfortranarray(_f64(B))
    return lib().ko_tr_prod_upper(A.shape[0], A.ctypes.data_as(_dp), B.ctypes.data_as(_dp))


def criterion_functional(tnim, is_inv, crit):
    A = np.asfortranarray(_f64(tnim)).copy(order="F")
    exc = C.c_int(0)
    v = lib().ko_criterion_functional(A.shape[0], A.ctypes.data_as(_dp), int(is_inv), crit, C.byref(exc))
    if exc.value:
        raise ArithmeticError("SingularException")
    return v


def apply_transformation(M, DT, r_is_inv=False):
    """Returns (t_x_t, t_is_inv).  DT None = Identity."""
    M = np.asfortranarray(_f64(M)).copy(order="F")
    r = M.shape[0]
    if DT is None:
        out = np.empty((r, r), order="F")
        flag = C.c_int(0)
        lib().ko_apply_transformation(r, r, TRAFO_IDENTITY, M.ctypes.data_as(_dp), int(r_is_inv), None,
                                      out.ctypes.data_as(_dp), C.byref(flag))
        return np.array(out), bool(flag.value)
    DT = np.asfortranarray(_f64(DT))
    t = DT.shape[0]
    out = np.empty((t, t), order="F")
    flag = C.c_int(0)
    rc = lib().ko_apply_transformation(r, t, TRAFO_DELTA, M.ctypes.data_as(_dp), int(r_is_inv),
                                       DT.ctypes.data_as(_dp), out.ctypes.data_as(_dp), C.byref(flag))
    if rc:
        raise ArithmeticError("PosDefException")
    return np.array(out), bool(flag.value)


def move_designpoint_how_far(dp, v, lb, ub):
    dp, v, lb, ub = map(_f64, (dp, v, lb, ub))
    return lib().ko_move_designpoint_how_far(len(dp), _d(dp), _d(v), _d(lb), _d(ub))


def move_handle_fixed(v_atoms, v_w, fixw, fixp):
    va = _f64(v_atoms).copy()
    vw = _f64(v_w).copy()
    K = len(vw)
    lib().ko_move_handle_fixed(K, va.size // K, _d(va), _d(vw), _i32(fixw).ctypes.data_as(_ip),
                               _i32(fixp).ctypes.data_as(_ip))
    return va, vw


def move_how_far(pts, w, v_atoms, v_w, lb, ub):
    pts, w, va, vw, lb, ub = map(_f64, (pts, w, v_atoms, v_w, lb, ub))
    K = len(w)
    return lib().ko_move_how_far(K, pts.size // K, _d(pts), _d(w), _d(va), _d(vw), _d(lb), _d(ub))


def move_add_v(pts, w, t, v_atoms, v_w, lb, ub, fixw):
    pts = _f64(pts).copy()
    w = _f64(w).copy()
    va, vw, lb, ub = map(_f64, (v_atoms, v_w, lb, ub))
    K = len(w)
    lib().ko_move_add_v(K, pts.size // K, _d(pts), _d(w), t, _d(va), _d(vw), _d(lb), _d(ub),
                        _i32(fixw).ctypes.data_as(_ip))
    return pts, w


def ap_add(pts, w, v_atoms, v_w, lb, ub, fixw, fixp):
    pts = _f64(pts).copy()
    w = _f64(w).copy()
    va = _f64(v_atoms).copy()
    vw = _f64(v_w).copy()
    lb, ub = _f64(lb), _f64(ub)
    K = len(w)
    t = lib().ko_ap_add(K, pts.size // K, _d(pts), _d(w), _d(va), _d(vw), _d(lb), _d(ub),
                        _i32(fixw).ctypes.data_as(_ip), _i32(fixp).ctypes.data_as(_ip))
    return pts, w, va, vw, t


def philox4x32(seed, c0, c1, c2, c3):
    out = (C.c_uint32 * 4)()
    lib().ko_philox4x32(seed, c0, c1, c2, c3, out)
    return list(out)


def uniform2(seed, it, particle, comp, stream):
    out = np.empty(2)
    lib().ko_uniform2_export(seed, it, particle, comp, stream, _d(out))
    return out


class Pso:
    """Oracle PSO (src/pso.jl + src/optimize.jl:42-70) on the shared Philox stream."""

    def __init__(self, problem, S, K, lb, ub, fixw=None, fixp=None, seed=1, c1=2.05, c2=2.05,
                 mode=0, gconst=None, nthreads=1):
        self.problem, self.S, self.K, self.d = problem, S, K, problem.d
        self._lb, self._ub = _f64(lb), _f64(ub)
        self._fixw = _i32(np.zeros(K) if fixw is None else fixw)
        self._fixp = _i32(np.zeros(K) if fixp is None else fixp)
        gA = gtrB = None
        if gconst is not None:
            self._gA = _f64(np.transpose(gconst[0], (0, 2, 1)))
            self._gtrB = _f64(gconst[1])
            gA, gtrB = _d(self._gA), _d(self._gtrB)
        self._h = lib().ko_pso_create(problem._h, S, K, c1, c2, _d(self._lb), _d(self._ub),
                                      self._fixw.ctypes.data_as(_ip), self._fixp.ctypes.data_as(_ip),
                                      seed, mode, gA, gtrB, nthreads)

    def init(self, proto_points, proto_weights):
        w = _f64(proto_weights).reshape(-1, self.K)
        pts = _f64(proto_points).reshape(len(w), self.K, self.d)
        rc = lib().ko_pso_init(self._h, _d(pts), _d(w), len(w))
        if rc == 1:
            raise RuntimeError("infinite objective value at best initial point")
        if rc == 2:
            raise RuntimeError("objective evaluates to NaN on all initial points")

    def tick(self):
        lib().ko_pso_tick(self._h)

    def state(self):
        S, K, d = self.S, self.K, self.d
        st = dict(x_pts=np.empty((S, K, d)), x_w=np.empty((S, K)), p_pts=np.empty((S, K, d)),
                  p_w=np.empty((S, K)), g_pts=np.empty((K, d)), g_w=np.empty(K),
                  v_pts=np.empty((S, K, d)), v_w=np.empty((S, K)), fx=np.empty(S), fp=np.empty(S))
        fg = C.c_double(0)
        ne = C.c_long(0)
        lib().ko_pso_get_state(self._h, _d(st["x_pts"]), _d(st["x_w"]), _d(st["p_pts"]), _d(st["p_w"]),
                               _d(st["g_pts"]), _d(st["g_w"]), _d(st["v_pts"]), _d(st["v_w"]),
                               _d(st["fx"]), _d(st["fp"]), C.cast(C.byref(fg), _dp), C.byref(ne))
        st["fg"], st["n_eval"] = fg.value, ne.value
        return st

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.ko_pso_destroy(self._h)
            self._h = None
