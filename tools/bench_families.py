"""Evaluation-kernel throughput of the families outside the BASELINE configs (vector units, dose-time-response,
logistic regression, dense covariance) at 1024 particles x 10^6 draws (2 x 10^5 for the dense family): tuning aid."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

F = kb._ffi


def run(name, h, pts, w):
    h.objective_batch(pts, w)
    h.profile_enable(True)
    for _ in range(3):
        fx = h.objective_batch(pts, w)
    ms, cnt = h.profile_read()
    S = len(w)
    print("%-22s kernel %8.3f ms  %.3e evals/s  finite %d/%d" % (name, ms / cnt, S * h.N / (ms / cnt * 1e-3), np.isfinite(fx).sum(), S),
          flush=True)
    h.close()


n, S = 1_000_000, 1024
pw = np.full(n, 1.0 / n)
th = wl.tpc_prior(n, seed=4)
pts, w = wl.jittered_swarm(S, 3, 0.0, 24.0, seed=1, log_lo=0.2)
run("VU m=3 K=3", kb.ProblemHandle(F.MODEL_VU, [0.1], th, pw, 0, None, [0.0], [24.0], m=3, cp_id=1), pts, w)
thd = wl.dtr_prior(n)
ptd, wd = wl.jittered_swarm_nd(S, 6, [0.0, 0.0], [100.0, 24.0], seed=2)
run("DTR CopyBoth K=6", kb.ProblemHandle(F.MODEL_DTR, [1.0], thd, pw, 0, None, [0.0, 0.0], [100.0, 24.0], m=1, cp_id=2), ptd, wd)
rng = np.random.default_rng(0)
thl = np.column_stack([rng.normal(-1.0, 0.5, n), rng.normal(1.5, 0.7, n), rng.normal(1.5, 0.7, n)])
ptl = rng.uniform(0, 1, (S, 6, 2))
wlr = rng.dirichlet(np.full(6, 4.0), S)
run("LinLogReg d=2 K=6", kb.ProblemHandle(F.MODEL_LINLOGREG, [0.0], thl, pw, 0, None, [0.0, 0.0], [1.0, 1.0]), ptl, wlr)
ths = np.column_stack([rng.uniform(0.5, 1.5, n), rng.uniform(1.0, 2.0, n), rng.uniform(0.0, 0.5, n)])
pts1, w1 = wl.jittered_swarm(S, 6, 0.0, 1.0, seed=3)
run("SinLogReg K=6", kb.ProblemHandle(F.MODEL_SINLOGREG, [0.0], ths, pw, 0, None, [0.0], [1.0]), pts1, w1)
n2 = 200_000
pts2, w2 = wl.jittered_swarm(S, 3, 0.0, 6.0, seed=1, log_lo=0.2)
run("VUAR m=3 K=3 (2e5)", kb.ProblemHandle(F.MODEL_VUAR, [0.1, 0.5], th[:n2], np.full(n2, 1.0 / n2), 0, None, [0.0], [6.0], m=3, cp_id=1), pts2, w2)
