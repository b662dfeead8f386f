import os, sys, time, numpy as np
sys.path.insert(0, '.')
import kirstine_b200 as kb
from tests import workloads as wl
def run(name, model, mp, theta, ub, G, m=1, cp=0):
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), 0, None, [0.0], [ub], m=m, cp_id=cp)
    at_p, at_w = np.array([2.0, 8.0, 20.0, 22.0]) * ub / 24.0, [0.3, 0.3, 0.2, 0.2]
    dirs = np.linspace(0.01, ub, G)
    h.gateaux_prepare(at_p, at_w)
    for _ in range(3): h.gateaux_eval(dirs)
    t0 = time.perf_counter()
    for _ in range(20): h.gateaux_eval(dirs)
    wall = (time.perf_counter() - t0) / 20
    print("%-6s G=%d N=%d: %.3f ms = %.3e /s" % (name, G, n, wall * 1e3, G * n / wall))
run("vu3", 4, [0.1], wl.tpc_prior(100000, seed=4), 24.0, 10000, m=3, cp=1)
run("tpc", 3, [1.0], wl.tpc_prior(100000, seed=4), 48.0, 10000)
run("sig", 2, [1.0], wl.sigemax_prior(100000), 1.0, 10000)
run("emax", 1, [1.0], wl.emax_prior(100000), 10.0, 10000)
