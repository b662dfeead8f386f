"""Time to solution on BASELINE config 5: sigmoid Emax, 10^6 prior draws, Pso(swarmsize = 1024, iterations = 500),
8-point equidistant prototype, through the public API (solve = DirectMaximization + simplify + sort)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 500
t0 = time.perf_counter()
dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 1)), model=kb.SigEmaxModel(sigma=1),
                      covariate_parameterization=kb.JustCopy("dose"),
                      prior_knowledge=kb.PriorSample(wl.sigemax_prior(n, seed=5)))
kb.seed(31415)
st = kb.DirectMaximization(optimizer=kb.Pso(iterations=iters, swarmsize=1024),
                           prototype=kb.equidistant_design(kb.region(dp), 8))
dp.handle().profile_enable(True)
t1 = time.perf_counter()
s, r = kb.solve(dp, st, maxweight=1e-3, maxdist=1e-2)
t2 = time.perf_counter()
kms, kcnt = dp.handle().profile_read()
dp.handle().profile_enable(False)
print('objective kernel: %d launches, %.3f ms average, %.2f s total' % (kcnt, kms / kcnt, kms / 1e3))
x, gd = kb.gateauxderivative_grid(s, dp)
t3 = time.perf_counter()
res = kb.optimization_result(r)
print("draws %d, iterations %d, n_eval %d" % (n, iters, res.n_eval))
print("problem set-up %.2f s, solve %.2f s (%.3e objective evaluations/s incl. simplify), Gateaux grid %.3f s"
      % (t1 - t0, t2 - t1, res.n_eval * n / (t2 - t1), t3 - t2))
print("objective %.6f, max Gateaux derivative on the grid %.2e" % (kb.objective(s, dp), gd.max()))
print(s)
