import sys, numpy as np
sys.path.insert(0, '.')
import kirstine_b200 as kb, oracle as ko
from tests import workloads as wl, referee as rf
theta = wl.sigemax_prior(1000)
pw = np.full(1000, 1e-3)
pts, w = wl.jittered_swarm(64, 8, 0.0, 1.0, seed=1)
h = kb.ProblemHandle(2, [1.0], theta, pw, 0, None, [0.0], [1.0])
got = h.objective_batch(pts, w)
want = ko.Problem(ko.SIGEMAX, [1.0], theta).objective_batch(pts, w, nthreads=8)
ref = np.array([rf.objective(2, [1.0], pts[i], w[i], theta, pw) for i in range(64)])
eg, eo = np.abs(got - ref) / np.abs(ref), np.abs(want - ref) / np.abs(ref)
print("engine vs referee: max %.2e median %.2e | oracle vs referee: max %.2e median %.2e | engine vs oracle max %.2e"
      % (eg.max(), np.median(eg), eo.max(), np.median(eo), (np.abs(got - want) / np.abs(want)).max()))
