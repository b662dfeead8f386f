"""Driver for ncu and for quick timings: launches every distinct hot kernel of the BASELINE configs once cold (the first
launch of each also loads its code) and then 5 times warm (`once`: one warm launch, for `ncu -k regex:... ` captures,
where every second captured launch is the warm one).
  cfg5  k_objective<KstSigEmax, D, Identity, uniform>   1024 particles x 1e6 draws (static swarm, all rows live)
  cfg2  k_objective<KstSigEmax, A, DeltaMethod t=1>     256 x 1e4  (and 1024 x 1e6 with `big`)
  cfg3  k_objective<KstTpc, D, Identity, uniform>       256 x 1e5
  cfg1  k_objective<KstEmax, D, Identity, uniform>      50 x 1e3   (and 1024 x 1e6 with `big`)
  cfg4  k_gateaux_eval<KstVu, 3>                        1e4 directions x 1e3 draws, and x 1e5 draws
Usage: python tools/profile_kernels.py [big] [once]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

big = "big" in sys.argv[1:]
REPS = 1 if "once" in sys.argv[1:] else 5   # `once`: one warm launch after the cold one (for ncu captures)


def objective(name, model, mp, theta, ub, K, S, crit=0, dt=None):
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), crit, dt, [0.0], [ub])
    pts, w = wl.random_swarm(S, K, 0.0, ub, seed=5)
    fx = h.objective_batch(pts, w)            # first launch: loads the kernel's code
    h.profile_enable(True)
    for _ in range(REPS):
        fx = h.objective_batch(pts, w)
    ms, cnt = h.profile_read()
    print("%-6s S=%5d N=%8d: kernel %9.4f ms (mean of %d), %.3e evals/s, fx[0] = %.12g" %
          (name, S, n, ms / cnt, cnt, S * n / (ms / cnt * 1e-3), fx[0]), flush=True)
    h.close()


def gateaux(name, theta, G):
    n = len(theta)
    h = kb.ProblemHandle(kb._ffi.MODEL_VU, [0.1], theta, np.full(n, 1.0 / n), 0, None, [0.0], [24.0], m=3,
                         cp_id=kb._ffi.CP_EQUITIME)
    assert not h.gateaux_prepare([2.0, 8.0, 20.0], [0.3, 0.3, 0.4])
    dirs = np.linspace(0.0, 24.0, G)
    gd = h.gateaux_eval(dirs)
    h.profile_enable(True)
    for _ in range(REPS):
        gd = h.gateaux_eval(dirs)
    ms, cnt = h.profile_read()
    print("%-6s G=%5d N=%8d: kernel %9.4f ms (mean of %d), %.3e (direction x draw)/s, max gd = %.6g" %
          (name, G, n, ms / cnt, cnt, G * n / (ms / cnt * 1e-3), gd.max()), flush=True)
    h.close()


def swarm(name, model, mp, theta, ub, K, S, iters):
    """A whole optimize() of a small problem: ONE launch of the persistent swarm kernel (k_pso_fused)."""
    import time
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), 0, None, [0.0], [ub])
    proto = (np.linspace(0.05 * ub, ub, K).reshape(K, 1), np.full(K, 1.0 / K))
    h.pso_optimize(S, K, iters, *proto, seed=1, state0=False)
    t0 = time.perf_counter()
    for _ in range(REPS):
        st, fx, *_ = h.pso_optimize(S, K, iters, *proto, seed=1, state0=False)
    dt = (time.perf_counter() - t0) / REPS
    print("%-6s S=%5d N=%8d: optimize(%d iterations) %8.3f ms wall = %.2f us per iteration, fg = %.10g" %
          (name, S, n, iters, dt * 1e3, dt * 1e6 / iters, fx[-1]), flush=True)
    h.close()


objective("cfg5", 2, [1.0], wl.sigemax_prior(1_000_000, seed=5), 1.0, 8, 1024)
if big:
    th = wl.sigemax_prior(1_000_000, seed=2)
    objective("cfg2L", 2, [1.0], th, 1.0, 6, 1024, crit=1, dt=wl.ded90_jacobians(th))
    objective("cfg3L", 3, [1.0], wl.tpc_prior(1_000_000), 48.0, 6, 1024)
    objective("cfg1L", 1, [1.0], wl.emax_prior(1_000_000), 10.0, 3, 1024)
    gateaux("cfg4L", wl.tpc_prior(100_000, seed=4), 10_000)
else:
    th = wl.sigemax_prior(10_000, seed=2)
    objective("cfg2", 2, [1.0], th, 1.0, 6, 256, crit=1, dt=wl.ded90_jacobians(th))
    objective("cfg3", 3, [1.0], wl.tpc_prior(100_000), 48.0, 6, 256)
    objective("cfg1", 1, [1.0], wl.emax_prior(1000), 10.0, 3, 50)
    gateaux("cfg4", wl.tpc_prior(1000, seed=4), 10_000)
    swarm("pso1", 1, [1.0], wl.emax_prior(1000), 10.0, 3, 50, 100)
    swarm("pso1b", 2, [1.0], wl.sigemax_prior(1000), 1.0, 8, 100, 50)
