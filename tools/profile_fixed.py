"""Where does a short CTA of k_objective spend its time?  1024 particles against just enough draws for one pair per
thread (37 x 256 x 2), so that the fixed per-CTA work dominates: driver for an `ncu --import-source on` capture whose
source page (stall samples per line) is read afterwards."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 37 * 256 * 2
theta = wl.sigemax_prior(n, seed=5)
pts, w = wl.random_swarm(1024, 8, 0.0, 1.0, seed=5)
h = kb.ProblemHandle(2, [1.0], theta, np.full(n, 1.0 / n), 0, None, [0.0], [1.0])
fx = h.objective_batch(pts, w)
h.profile_enable(True)
for _ in range(10):
    fx = h.objective_batch(pts, w)
ms, cnt = h.profile_read()
print("N = %d: kernel %.4f ms" % (n, ms / cnt))
