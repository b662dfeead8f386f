"""Small driver for ncu: evaluates the cfg5 swarm (static, all K = 8 rows live) a few times through
kst_objective_batch.  Usage: python tools/profile_objective.py [reps] [draws] [swarm] [cfg]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
S = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
cfg = sys.argv[4] if len(sys.argv) > 4 else "5"
if cfg == "5":
    theta, model, mp, ub, K, crit, dt = wl.sigemax_prior(n, seed=5), 2, [1.0], 1.0, 8, 0, None
elif cfg == "2":
    theta = wl.sigemax_prior(n, seed=2)
    model, mp, ub, K, crit, dt = 2, [1.0], 1.0, 6, 1, wl.ded90_jacobians(theta)
elif cfg == "3":
    theta, model, mp, ub, K, crit, dt = wl.tpc_prior(n), 3, [1.0], 48.0, 6, 0, None
elif cfg == "1":
    theta, model, mp, ub, K, crit, dt = wl.emax_prior(n), 1, [1.0], 10.0, 3, 0, None
pts, w = wl.random_swarm(S, K, 0.0, ub, seed=5)
h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), crit, dt, [0.0], [ub])
h.profile_enable(True)
for _ in range(reps):
    fx = h.objective_batch(pts, w)
ms, cnt = h.profile_read()
print("cfg", cfg, "kernel ms avg", ms / cnt, "evals/s", S * n / (ms / cnt * 1e-3), "fx[0]", fx[0])
