"""Ad-hoc hunt: random PSO set-ups (family, K, swarm size, fixed weights / points, number of prototypes, region) for a
few ticks, engine state against the oracle's on the shared Philox stream.  Divergence caused by a rounding-level
difference in an objective value flipping a best-update is reported separately from a mismatch in the move itself
(positions are compared only while the best-update decisions agree)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb  # noqa: E402
import oracle as ko  # noqa: E402
from tests import workloads as wl  # noqa: E402

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
bad = flips = 0
for seed in range(n_cases):
    rng = np.random.default_rng(seed)
    fam = seed % 3
    if fam == 0:
        model, mp, th, lb, ub, d, okw = 1, [1.0], wl.emax_prior(100, seed=seed), [0.0], [10.0], 1, {}
    elif fam == 1:
        model, mp, th, lb, ub, d, okw = 2, [1.0], wl.sigemax_prior(100, seed=seed), [0.0], [1.0], 1, {}
    else:
        model, mp, th, lb, ub, d = 5, [1.0], wl.dtr_prior(60, seed=seed), [0.0, 0.0], [100.0, 24.0], 2
        okw = dict(m=1, d=2, cp_id=2)
    r = th.shape[1]
    K = int(rng.integers(r, r + 4))
    S = int(rng.choice([2, 5, 16, 33]))
    P = int(rng.integers(1, min(S, 3) + 1))
    fixw = (rng.random(K) < 0.25).astype(int)
    fixp = (rng.random(K) < 0.25).astype(int)
    pw = np.full(len(th), 1.0 / len(th))
    if d == 1:
        pp, ww = wl.jittered_swarm(P, K, lb[0], ub[0], seed=seed)
    else:
        pp, ww = wl.jittered_swarm_nd(P, K, lb, ub, seed=seed)
    h = kb.ProblemHandle(model, mp, th, pw, 0, None, lb, ub, **({"m": 1, "cp_id": 2} if fam == 2 else {}))
    o = ko.Problem(model, mp, th, pw, **okw)
    if h.pso_init(S, K, pp, ww, fixw=fixw, fixp=fixp, seed=seed + 1) != 0:
        continue
    po = ko.Pso(o, S, K, lb, ub, fixw=fixw, fixp=fixp, seed=seed + 1)
    po.init(pp, ww)
    ok = True
    for it in range(6):
        a, b = h.pso_state(), po.state()
        same_best = np.allclose(a["p_pts"], b["p_pts"], rtol=1e-9, atol=1e-12) and np.allclose(a["g_pts"], b["g_pts"], rtol=1e-9, atol=1e-12)
        if not same_best:
            flips += 1
            break
        for key in ("x_pts", "x_w", "v_pts", "v_w"):
            if not np.allclose(a[key], b[key], rtol=1e-10, atol=1e-13):
                print("MISMATCH seed", seed, "fam", fam, "K", K, "S", S, "P", P, "fixw", fixw, "fixp", fixp, "tick", it, key,
                      np.abs(a[key] - b[key]).max())
                ok = False
                break
        if not ok:
            break
        h.pso_run(1)
        po.tick()
    bad += not ok
print("cases", n_cases, "mismatches", bad, "runs ended early by a flipped best-update", flips)
