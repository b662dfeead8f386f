"""Latency of the Exchange strategy on BASELINE config 4 (vector-response model, 1000 draws, Pso(25, 10) x 2, 8 steps)
and of a small DirectMaximization (config 1), through the public API."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
from tests import workloads as wl  # noqa: E402

dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(delta=(0, 24)), model=kb.VUMod(0.1, 3),
                      covariate_parameterization=kb.EquiTime(), prior_knowledge=kb.PriorSample(wl.tpc_prior(1000, seed=4)))
st = kb.Exchange(optimizer_direction=kb.Pso(swarmsize=25, iterations=10), optimizer_weight=kb.Pso(swarmsize=25, iterations=10),
                 steps=8, candidate=kb.uniform_design([[2.0], [8.0], [20.0]]))
kb.seed(1)
kb.solve(dp, st)          # warm-up (context, module load)
t0 = time.perf_counter()
s, r = kb.solve(dp, st)
t1 = time.perf_counter()
print("Exchange, 8 steps: %.2f ms" % ((t1 - t0) * 1e3))
dp1 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 10)), model=kb.EmaxModel(1),
                       covariate_parameterization=kb.CopyDose(), prior_knowledge=kb.PriorSample(wl.emax_prior(1000)))
st1 = kb.DirectMaximization(optimizer=kb.Pso(swarmsize=50, iterations=100), prototype=kb.equidistant_design(kb.region(dp1), 3))
kb.solve(dp1, st1)
t0 = time.perf_counter()
kb.solve(dp1, st1)
t1 = time.perf_counter()
print("DirectMaximization cfg 1 (50 x 100 x 1000): %.2f ms" % ((t1 - t0) * 1e3))
pr = cProfile.Profile()
pr.enable()
kb.solve(dp, st)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
