"""Throughput of every BASELINE.json config (tuning aid; the contract benchmark is bench.py).
cfg 1, 1b, 2, 3: objective evals/s through kst_objective_batch on a static swarm and PSO iterations/s;
cfg 4: Gateaux grid evals/s (directions x draws)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402


def run_objective(name, model, mp, theta, ub, K, S, crit=0, dt=None, iters=100, m=1, cp=0):
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), crit, dt, [0.0], [ub], m=m, cp_id=cp)
    pts, w = wl.random_swarm(S, K, 0.0, ub, seed=5)
    for _ in range(3):
        h.objective_batch(pts, w)
    h.profile_enable(True)
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        h.objective_batch(pts, w)
    wall = (time.perf_counter() - t0) / reps
    ms, cnt = h.profile_read()
    h.profile_enable(False)
    st = h.pso_init(S, K, pts[0], w[0], seed=1)
    h.pso_run(3)
    t0 = time.perf_counter()
    h.pso_run(iters)
    pso = (time.perf_counter() - t0) / iters
    print("%-4s S=%5d N=%8d K=%d: kernel %8.3f ms  e2e call %8.3f ms  %.3e evals/s (kernel)  PSO iter %8.3f ms = %.3e evals/s"
          % (name, S, n, K, ms / cnt, wall * 1e3, S * n / (ms / cnt * 1e-3), pso * 1e3, S * n / pso), flush=True)


def run_gateaux(name, model, mp, theta, ub, G, m=1, cp=0):
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), 0, None, [0.0], [ub], m=m, cp_id=cp)
    at_p, at_w = np.array([2.0, 8.0, 20.0]) * ub / 24.0, [0.3, 0.3, 0.4]
    dirs = np.linspace(0, ub, G)
    t0 = time.perf_counter()
    h.gateaux_prepare(at_p, at_w)
    prep = time.perf_counter() - t0
    for _ in range(3):
        h.gateaux_eval(dirs)
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        h.gateaux_eval(dirs)
    wall = (time.perf_counter() - t0) / reps
    print("%-4s G=%6d N=%8d: prepare %7.3f ms  eval call %8.3f ms = %.3e (direction x draw)/s" %
          (name, G, n, prep * 1e3, wall * 1e3, G * n / wall), flush=True)


if __name__ == "__main__":
    run_objective("1", 1, [1.0], wl.emax_prior(1000), 10.0, 3, 50)
    run_objective("1b", 2, [1.0], wl.sigemax_prior(1000), 1.0, 8, 100)
    th = wl.sigemax_prior(10_000, seed=2)
    run_objective("2", 2, [1.0], th, 1.0, 6, 256, crit=1, dt=wl.ded90_jacobians(th))
    run_objective("3", 3, [1.0], wl.tpc_prior(100_000), 48.0, 6, 256)
    run_gateaux("4", 4, [0.1], wl.tpc_prior(1000, seed=4), 24.0, 10_000, m=3, cp=1)
    run_gateaux("4L", 4, [0.1], wl.tpc_prior(100_000, seed=4), 24.0, 10_000, m=3, cp=1)
    run_objective("5", 2, [1.0], wl.sigemax_prior(1_000_000, seed=5), 1.0, 8, 1024, iters=10)
