"""Latency of the Exchange strategy on BASELINE config 4 (vector-response model, 1000 draws, Pso(25, 10) x 2, 8 steps):
the C-ABI call alone (kst_exchange_run), the whole solve() through the public API, and the host-stepped loop
(trace_state = True) for comparison; plus a small DirectMaximization (config 1)."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402


def best_of(fn, n=5):
    best = None
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best * 1e3


dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(delta=(0, 24)), model=kb.VUMod(0.1, 3),
                      covariate_parameterization=kb.EquiTime(), prior_knowledge=kb.PriorSample(wl.tpc_prior(1000, seed=4)))
st = kb.Exchange(optimizer_direction=kb.Pso(swarmsize=25, iterations=10), optimizer_weight=kb.Pso(swarmsize=25, iterations=10),
                 steps=8, candidate=kb.uniform_design([[2.0], [8.0], [20.0]]))
kb.seed(1)
kb.solve(dp, st)          # warm-up (context, module load)
h = dp.handle()
cand = st.candidate
print("kst_exchange_run alone, 8 steps, with initial-state snapshots: %.2f ms" %
      best_of(lambda: h.exchange_run(cand.points, cand.weights, 8, (25, 10, 2.05, 2.05), (25, 10, 2.05, 2.05), seed=1)))
print("kst_exchange_run alone, 8 steps, traces only:                  %.2f ms" %
      best_of(lambda: h.exchange_run(cand.points, cand.weights, 8, (25, 10, 2.05, 2.05), (25, 10, 2.05, 2.05), seed=1, state0=False)))
print("solve(dp, Exchange) through the public API:                    %.2f ms" % best_of(lambda: kb.solve(dp, st)))
os.environ["KST_NO_FUSED"] = "1"
print("  the same with three launches per iteration (KST_NO_FUSED):   %.2f ms" % best_of(lambda: kb.solve(dp, st)))
del os.environ["KST_NO_FUSED"]
print("solve(dp, Exchange, trace_state=True), host-stepped loop:      %.2f ms" % best_of(lambda: kb.solve(dp, st, trace_state=True), 3))
dp1 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 10)), model=kb.EmaxModel(1),
                       covariate_parameterization=kb.CopyDose(), prior_knowledge=kb.PriorSample(wl.emax_prior(1000)))
st1 = kb.DirectMaximization(optimizer=kb.Pso(swarmsize=50, iterations=100), prototype=kb.equidistant_design(kb.region(dp1), 3))
kb.solve(dp1, st1)
print("DirectMaximization cfg 1 (50 x 100 x 1000):                    %.2f ms" % best_of(lambda: kb.solve(dp1, st1)))
if "--profile" in sys.argv:
    pr = cProfile.Profile()
    pr.enable()
    kb.solve(dp, st)
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
