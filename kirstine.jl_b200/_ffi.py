"""ctypes binding of libkirstine_b200.so (include/kirstine_b200.h).

There is no CPU fallback: loading fails loudly when the library has not been built, and every
compute call fails loudly when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# KST_LIB_PATH selects an alternative build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("KST_LIB_PATH") or os.path.join(_HERE, "libkirstine_b200.so")

MODEL_EMAX, MODEL_SIGEMAX, MODEL_TPC, MODEL_VU, MODEL_DTR, MODEL_LINLOGREG, MODEL_SINLOGREG = 1, 2, 3, 4, 5, 6, 7
MODEL_VUAR = 8
CP_COPY, CP_EQUITIME = 0, 1
CP_DTR_COPYBOTH, CP_DTR_FIXEDDOSE, CP_DTR_FIXEDTIMES, CP_DTR_EQUIDISTANT, CP_DTR_LOGEQUIDISTANT = 2, 3, 4, 5, 6
CRIT_D, CRIT_A, CRIT_DEXP = 0, 1, 2
TRAFO_IDENTITY, TRAFO_DELTA = 0, 1
PSO_MODE_OBJECTIVE, PSO_MODE_GATEAUX = 0, 1

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p

# name -> (restype, argtypes); every symbol include/kirstine_b200.h declares
SIGNATURES = {
    "kst_last_error": (C.c_char_p, []),
    "kst_version": (C.c_int, []),
    "kst_device_count": (C.c_int, [_ip]),
    "kst_problem_create": (C.c_int, [C.c_int, _dp, C.c_int, C.c_int, _dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_int, C.POINTER(_vp)]),
    "kst_problem_destroy": (C.c_int, [_vp]),
    "kst_comm_unique_id": (C.c_int, [_vp]),
    "kst_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "kst_comm_peer_handle": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "kst_comm_peer_open": (C.c_int, [_vp, _vp]),
    "kst_comm_peer_error": (C.c_int, [_vp, _ip]),
    "kst_comm_peer_close": (C.c_int, [_vp]),
    "kst_objective_batch": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp]),
    "kst_objective_batch_device": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp]),
    "kst_objective_batch_noalloc": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp]),
    "kst_stream_synchronize": (C.c_int, [_vp]),
    "kst_stream": (_vp, [_vp]),
    "kst_problem_set_regularization": (C.c_int, [_vp, _dp]),
    "kst_informationmatrix": (C.c_int, [_vp, C.c_int, _dp, _dp, C.c_int64, _dp]),
    "kst_gateaux_prepare": (C.c_int, [_vp, C.c_int, _dp, _dp, _ip]),
    "kst_gateaux_constants": (C.c_int, [_vp, _dp, _dp]),
    "kst_gateaux_eval": (C.c_int, [_vp, C.c_int64, _dp, _dp]),
    "kst_pso_init": (C.c_int, [_vp, C.c_int, C.c_int, C.c_double, C.c_double, _dp, _dp, C.c_int, _ip, _ip,
                               C.c_uint64, C.c_int, _ip]),
    "kst_pso_run": (C.c_int, [_vp, C.c_int, _dp, _dp, _dp]),
    "kst_pso_optimize": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _dp, _dp, C.c_int, _ip, _ip,
                                   C.c_uint64, C.c_int, _ip, _dp, _dp, _dp, _dp]),
    "kst_pso_get_state": (C.c_int, [_vp] + [_dp] * 11 + [C.POINTER(C.c_int64)]),
    "kst_profile_enable": (C.c_int, [_vp, C.c_int]),
    "kst_profile_read": (C.c_int, [_vp, _dp, C.POINTER(C.c_int64)]),
    "kst_fp64_peak_probe": (C.c_int, [C.c_int, _dp]),
    "kst_math_selftest": (C.c_int, [C.c_int, C.c_int64, _dp, _dp, _dp, _dp]),
}



class ExchangeOptions(C.Structure):
    """kst_exchange_options (include/kirstine_b200.h)."""
    _fields_ = [("steps", C.c_int), ("swarm_direction", C.c_int), ("iters_direction", C.c_int),
                ("c1_direction", C.c_double), ("c2_direction", C.c_double), ("swarm_weight", C.c_int),
                ("iters_weight", C.c_int), ("c1_weight", C.c_double), ("c2_weight", C.c_double),
                ("maxweight", C.c_double), ("maxdist", C.c_double), ("unique_a", C.c_double), ("unique_b", C.c_double),
                ("seed", C.c_uint64)]


SIGNATURES["kst_exchange_run"] = (C.c_int, [_vp, C.POINTER(ExchangeOptions), C.c_int, _dp, _dp, C.c_int, _ip, _dp, _dp,
                                            _dp, _dp, _dp, _dp, _dp, _ip, _dp, _dp, _ip])

_lib = None


class EngineError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("kirstine_b200 error %d: %s" % (code, msg))
        self.code = code


def lib():
    """Load the shared library (no CUDA call is made by loading)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "libkirstine_b200.so is not built (%s); run `python __graft_entry__.py build` -- "
                "this engine has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise EngineError(rc, lib().kst_last_error().decode())


def f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def ptr(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def device_count():
    n = C.c_int(0)
    check(lib().kst_device_count(C.byref(n)))
    return n.value


class ProblemHandle:
    """Owns one kst_problem (one GPU, one shard of the prior sample).

    theta (N, r); prior_weights (N,); dt (N, t, r) or None.  Point arrays are (..., K, d).
    """

    def __init__(self, model_id, model_params, theta, prior_weights, crit, dt, lb, ub, m=1, cp_id=CP_COPY,
                 device=0, cp_params=()):
        theta = f64(theta)
        if theta.ndim != 2:
            raise ValueError("theta must be (N, r)")
        self.N, self.r = theta.shape
        pw = f64(prior_weights)
        lb, ub = f64(lb).ravel(), f64(ub).ravel()
        self.d = len(lb)
        self.m = int(m)
        self.crit = crit
        if dt is None:
            self.trafo, self.t, dtc = TRAFO_IDENTITY, self.r, None
        else:
            dt = f64(dt)
            self.trafo, self.t = TRAFO_DELTA, dt.shape[1]
            dtc = f64(np.transpose(dt, (0, 2, 1)))  # per draw column-major t x r
        mp = f64(model_params).ravel()
        cpp = f64(cp_params).ravel() if len(cp_params) else None
        h = _vp()
        check(lib().kst_problem_create(model_id, ptr(mp), len(mp), cp_id, ptr(cpp), 0 if cpp is None else len(cpp), crit,
                                       self.trafo, self.r, self.t, self.m,
                                       self.d, self.N, ptr(theta), ptr(pw), ptr(dtc), ptr(lb), ptr(ub), device,
                                       C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib().kst_problem_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- multi-GPU ------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * 128)()
        check(lib().kst_comm_unique_id(C.cast(buf, _vp)))
        return bytes(buf)

    def comm_init(self, nranks, rank, unique_id):
        buf = C.create_string_buffer(unique_id, 128)
        check(lib().kst_comm_init(self._h, nranks, rank, C.cast(buf, _vp)))

    def comm_peer_handle(self, nranks, rank):
        buf = (C.c_char * 64)()
        check(lib().kst_comm_peer_handle(self._h, nranks, rank, C.cast(buf, _vp)))
        return bytes(buf)

    def comm_peer_open(self, handles):
        """handles: list of the nranks 64-byte IPC handles in rank order."""
        blob = b"".join(handles)
        buf = C.create_string_buffer(blob, len(blob))
        check(lib().kst_comm_peer_open(self._h, C.cast(buf, _vp)))

    def comm_peer_close(self):
        check(lib().kst_comm_peer_close(self._h))

    def comm_peer_error(self):
        e = C.c_int(0)
        check(lib().kst_comm_peer_error(self._h, C.byref(e)))
        return e.value

    # -- objective --------------------------------------------------------------------------
    def objective_batch(self, points, weights):
        w = f64(weights)
        w = w.reshape(-1, w.shape[-1])
        S, K = w.shape
        pts = f64(points).reshape(S, K, self.d)
        fx = np.empty(S)
        check(lib().kst_objective_batch(self._h, S, K, pts.ctypes.data, w.ctypes.data, fx.ctypes.data))
        return fx

    def objective_batch_raw(self, S, K, points_ptr, weights_ptr, fx_ptr):
        """Host pointers as integers (e.g. pinned torch tensors' data_ptr())."""
        check(lib().kst_objective_batch(self._h, S, K, points_ptr, weights_ptr, fx_ptr))

    def objective_batch_device(self, S, K, d_points, d_weights, d_fx):
        check(lib().kst_objective_batch_device(self._h, S, K, d_points, d_weights, d_fx))

    def objective_batch_noalloc(self, S, K, d_points, d_weights, d_fx):
        """Device pointers (integers); evaluates without touching the CUDA allocator (fixed 2 MiB scratch
        inside the library, at most 4096 particles).  Asynchronous: call synchronize() before reading d_fx."""
        check(lib().kst_objective_batch_noalloc(self._h, S, K, d_points, d_weights, d_fx))

    def synchronize(self):
        check(lib().kst_stream_synchronize(self._h))

    def stream(self):
        return lib().kst_stream(self._h)

    def set_regularization(self, R):
        """RegularizedFisherMatrix(R) of docs/src/extend-approximation.md:69-79; None switches back to FisherMatrix."""
        if R is None:
            check(lib().kst_problem_set_regularization(self._h, None))
            return self
        R = f64(R)
        if R.shape != (self.r, self.r):
            raise ValueError("R must be (%d, %d)" % (self.r, self.r))
        Rc = np.ascontiguousarray(R.T)   # column-major
        check(lib().kst_problem_set_regularization(self._h, ptr(Rc)))
        return self

    def informationmatrix(self, points, weights, draw=0):
        w = f64(weights).ravel()
        pts = f64(points).reshape(len(w), self.d)
        M = np.empty((self.r, self.r))
        check(lib().kst_informationmatrix(self._h, len(w), ptr(pts), ptr(w), draw, ptr(M)))
        return M

    # -- Gateaux ----------------------------------------------------------------------------
    def gateaux_prepare(self, points, weights):
        w = f64(weights).ravel()
        pts = f64(points).reshape(len(w), self.d)
        s = C.c_int(0)
        check(lib().kst_gateaux_prepare(self._h, len(w), ptr(pts), ptr(w), C.byref(s)))
        return bool(s.value)

    def gateaux_constants(self):
        A = np.empty((self.N, self.r, self.r))
        trB = np.empty(self.N)
        check(lib().kst_gateaux_constants(self._h, ptr(A), ptr(trB)))
        return np.transpose(A, (0, 2, 1)).copy(), trB

    def gateaux_eval(self, directions):
        dirs = f64(directions).reshape(-1, self.d)
        out = np.empty(len(dirs))
        check(lib().kst_gateaux_eval(self._h, len(dirs), ptr(dirs), ptr(out)))
        return out

    # -- PSO ----------------------------------------------------------------------------------
    def pso_init(self, S, K, proto_points, proto_weights, fixw=None, fixp=None, seed=1, c1=2.05, c2=2.05,
                 mode=PSO_MODE_OBJECTIVE):
        pw = f64(proto_weights).reshape(-1, K)
        P = len(pw)
        pp = f64(proto_points).reshape(P, K, self.d)
        fw = i32(np.zeros(K) if fixw is None else fixw)
        fp = i32(np.zeros(K) if fixp is None else fixp)
        st = C.c_int(0)
        check(lib().kst_pso_init(self._h, S, K, c1, c2, ptr(pp), ptr(pw), P, fw.ctypes.data_as(_ip),
                                 fp.ctypes.data_as(_ip), seed, mode, C.byref(st)))
        self._pso_dims = (S, K)
        return st.value

    def pso_optimize(self, S, K, iterations, proto_points, proto_weights, fixw=None, fixp=None, seed=1, c1=2.05, c2=2.05,
                     mode=PSO_MODE_OBJECTIVE, state0=True):
        """optimize() in one call: (status, trace_fx[iterations], trace_points, trace_weights, state0 dict or None)."""
        pw = f64(proto_weights).reshape(-1, K)
        P = len(pw)
        pp = f64(proto_points).reshape(P, K, self.d)
        fw = i32(np.zeros(K) if fixw is None else fixw)
        fp = i32(np.zeros(K) if fixp is None else fixp)
        d = self.d
        tfx, tp, tw = np.empty(iterations), np.empty((iterations, K, d)), np.empty((iterations, K))
        snap = np.empty(3 * S * K * (d + 1) + K * (d + 1) + 2 * S + 1) if state0 else None
        st = C.c_int(0)
        check(lib().kst_pso_optimize(self._h, S, K, iterations, c1, c2, ptr(pp), ptr(pw), P, fw.ctypes.data_as(_ip),
                                     fp.ctypes.data_as(_ip), seed, mode, C.byref(st), ptr(tfx), ptr(tp), ptr(tw), ptr(snap)))
        self._pso_dims = (S, K)
        return st.value, tfx, tp, tw, (unpack_state(snap, S, K, d) if state0 else None)

    def exchange_run(self, cand_points, cand_weights, steps, opt_direction, opt_weight, maxweight=0.0, maxdist=0.0,
                     unique_a=0.0, unique_b=0.0, seed=1, state0=True):
        """kst_exchange_run.  opt_* = (swarmsize, iterations, c1, c2).  Returns a dict with the final candidate
        (points, weights) and per step the direction / weight traces (and initial states when state0)."""
        w0 = f64(cand_weights).ravel()
        K0, d = len(w0), self.d
        p0 = f64(cand_points).reshape(K0, d)
        Kmax = K0 + steps
        Sd, itd, Sw, itw = opt_direction[0], opt_direction[1], opt_weight[0], opt_weight[1]
        o = ExchangeOptions(steps, Sd, itd, opt_direction[2], opt_direction[3], Sw, itw, opt_weight[2], opt_weight[3],
                            maxweight, maxdist, unique_a, unique_b, seed)
        Kout, done = C.c_int(0), C.c_int(0)
        pts, w = np.empty((Kmax, d)), np.empty(Kmax)
        dfx, dx = np.empty((steps, itd)), np.empty((steps, itd, d))
        wfx, wp, ww = np.empty((steps, itw)), np.empty((steps, itw * Kmax * d)), np.empty((steps, itw * Kmax))
        wK = np.zeros(steps, dtype=np.int32)
        nsd = 3 * Sd * (d + 1) + (d + 1) + 2 * Sd + 1
        nsw = 3 * Sw * Kmax * (d + 1) + Kmax * (d + 1) + 2 * Sw + 1
        sd = np.empty((steps, nsd)) if state0 else None
        sw = np.empty((steps, nsw)) if state0 else None
        check(lib().kst_exchange_run(self._h, C.byref(o), K0, ptr(p0), ptr(w0), Kmax, C.byref(Kout), ptr(pts), ptr(w),
                                     ptr(dfx), ptr(dx), ptr(wfx), ptr(wp), ptr(ww), wK.ctypes.data_as(_ip), ptr(sd), ptr(sw),
                                     C.byref(done)))
        K = Kout.value
        out = dict(points=pts[:K].copy(), weights=w[:K].copy(), dir_fx=dfx, dir_x=dx, w_fx=wfx, w_K=wK, w_points=[],
                   w_weights=[], dir_state0=[], w_state0=[])
        for s_ in range(steps):
            k = int(wK[s_])
            out["w_points"].append(wp[s_, :itw * k * d].reshape(itw, k, d).copy())
            out["w_weights"].append(ww[s_, :itw * k].reshape(itw, k).copy())
            if state0:
                out["dir_state0"].append(unpack_state(sd[s_], Sd, 1, d))
                out["w_state0"].append(unpack_state(sw[s_], Sw, k, d))
        return out

    def pso_run_async(self, n_iters):
        """Enqueue n_iters swarm iterations without copying traces back (no host synchronisation)."""
        check(lib().kst_pso_run(self._h, n_iters, None, None, None))

    def pso_run(self, n_iters):
        S, K = self._pso_dims
        tfx = np.empty(n_iters)
        tp = np.empty((n_iters, K, self.d))
        tw = np.empty((n_iters, K))
        check(lib().kst_pso_run(self._h, n_iters, ptr(tfx), ptr(tp), ptr(tw)))
        return tfx, tp, tw

    def pso_state(self):
        S, K = self._pso_dims
        d = self.d
        st = dict(x_pts=np.empty((S, K, d)), x_w=np.empty((S, K)), p_pts=np.empty((S, K, d)), p_w=np.empty((S, K)),
                  g_pts=np.empty((K, d)), g_w=np.empty(K), v_pts=np.empty((S, K, d)), v_w=np.empty((S, K)),
                  fx=np.empty(S), fp=np.empty(S))
        fg = np.empty(1)
        ne = C.c_int64(0)
        check(lib().kst_pso_get_state(self._h, ptr(st["x_pts"]), ptr(st["x_w"]), ptr(st["p_pts"]), ptr(st["p_w"]),
                                      ptr(st["g_pts"]), ptr(st["g_w"]), ptr(st["v_pts"]), ptr(st["v_w"]),
                                      ptr(st["fx"]), ptr(st["fp"]), ptr(fg), C.byref(ne)))
        st["fg"], st["n_eval"] = float(fg[0]), ne.value
        return st

    # -- measurement --------------------------------------------------------------------------
    def profile_enable(self, on=True):
        """CUDA events around one stage's launches (kst_profile_enable): 1 / True = evaluation kernel, 2 = swarm move,
        3 = best update (several GPUs: the tail kernel with the peer exchange), 0 / False = off."""
        check(lib().kst_profile_enable(self._h, int(on)))

    def profile_read(self):
        ms = C.c_double(0)
        n = C.c_int64(0)
        check(lib().kst_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value


def unpack_state(snap, S, K, d):
    """The packed PsoState of kst_pso_optimize / kst_exchange_run as the dict pso_state() returns (n_eval = S)."""
    out, o = {}, 0
    for key, shape in (("x_pts", (S, K, d)), ("x_w", (S, K)), ("p_pts", (S, K, d)), ("p_w", (S, K)), ("v_pts", (S, K, d)),
                       ("v_w", (S, K)), ("g_pts", (K, d)), ("g_w", (K,)), ("fx", (S,)), ("fp", (S,))):
        n = int(np.prod(shape))
        out[key] = snap[o:o + n].reshape(shape).copy()
        o += n
    out["fg"], out["n_eval"] = float(snap[o]), S
    return out


def fp64_peak_probe(device=0):
    tf = C.c_double(0)
    check(lib().kst_fp64_peak_probe(device, C.byref(tf)))
    return tf.value


def math_selftest(x, device=0):
    x = f64(x).ravel()
    e, r, l = np.empty_like(x), np.empty_like(x), np.empty_like(x)
    check(lib().kst_math_selftest(device, len(x), ptr(x), ptr(e), ptr(r), ptr(l)))
    return e, r, l
