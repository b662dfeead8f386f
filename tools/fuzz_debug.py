import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kirstine_b200 as kb
import oracle as ko
from tests import referee as rf, workloads as wl
from tests.test_gpu_fuzz import FAMILIES, _orthonormal_rows
for seed in [int(a) for a in sys.argv[1:]]:
    rng = np.random.default_rng(1000 + seed)
    name = list(FAMILIES)[seed % 3]
    model, mp, prior, ub, log_lo, r = FAMILIES[name]
    N = int(rng.choice([1, 2, 31, 33, 255, 257, 700])); S = int(rng.choice([1, 2, 7, 8, 9, 15, 17, 33]))
    K = int(rng.integers(r, r + 9)); crit = int(rng.integers(0, 3)); t = int(rng.integers(0, r + 1))
    th = prior(N, seed=seed + 5)
    pw = rng.dirichlet(np.ones(N)) if rng.random() < 0.5 else np.full(N, 1.0 / N)
    dt = _orthonormal_rows(N, t, r, seed) if t else None
    pts, w = wl.jittered_swarm(S, K, 0.0, ub, seed=seed, log_lo=log_lo)
    for s in range(S):
        extra = K - r; nz = int(rng.integers(0, extra + 1)); idx = rng.choice(K, size=nz, replace=False)
        w[s, idx] = 0.0; w[s] /= w[s].sum()
        if extra - nz > 0 and rng.random() < 0.5:
            live = np.flatnonzero(w[s] > 0); pts[s, rng.choice(live), 0] = 0.0
    h = kb.ProblemHandle(model, mp, th, pw, crit, dt, (0.0,), (ub,)); o = ko.Problem(model, mp, th, pw, crit=crit, dt=dt)
    got, want = h.objective_batch(pts, w), o.objective_batch(pts, w)
    uniform = np.all(pw == pw[0])
    print("seed", seed, name, "N", N, "S", S, "K", K, "crit", crit, "t", t, "uniform_w", uniform)
    for i in range(S):
        if np.isfinite(want[i]) and abs(got[i] - want[i]) > 1e-10 * abs(want[i]):
            ref = rf.objective(model, mp, pts[i], w[i], th, pw, crit=crit, dt=dt)
            # conditioning of the worst draw
            conds = []
            for j in range(min(N, 50)):
                M = o.informationmatrix(pts[i], w[i], j); conds.append(np.linalg.cond(M))
            print("  particle", i, "got", got[i], "oracle", want[i], "referee", float(ref), "rel err engine %.1e oracle %.1e" % (abs(got[i]-ref)/abs(ref), abs(want[i]-ref)/abs(ref)),
                  "live rows", int((w[i] > 0).sum()), "zeros", int((pts[i,:,0]==0).sum()), "max cond %.1e" % max(conds))
    fin = np.isfinite(want) & np.isfinite(got)
    if fin.any():
        s0 = int(np.flatnonzero(fin)[0])
        dirs = np.linspace(ub * 0.01, ub, 37)
        if not h.gateaux_prepare(pts[s0], w[s0]):
            gd, wg = h.gateaux_eval(dirs), o.gateauxderivative(pts[s0], w[s0], dirs)
            ref = np.array([float(rf.gateaux(model, mp, pts[s0], w[s0], [z], th, pw, crit=crit, dt=dt)[0]) for z in dirs])
            sc = np.abs(ref).max()
            conds = [np.linalg.cond(o.informationmatrix(pts[s0], w[s0], j)) for j in range(min(N, 50))]
            print("  gateaux at particle", s0, "max err engine %.1e oracle %.1e (relative to max |gd| %.3g)  max cond %.1e  live rows %d"
                  % (np.abs(gd - ref).max() / sc, np.abs(wg - ref).max() / sc, sc, max(conds), int((w[s0] > 0).sum())))
