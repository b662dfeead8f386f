"""Regenerates tests/golden/oracle_fixtures.npz: seeded inputs and the CPU oracle's outputs for the five
BASELINE.json configs (at oracle-sized N), so that GPU parity can also be checked against committed numbers
and drift of the oracle itself is caught.  Run from the repo root: python tests/golden/make_fixtures.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as ko  # noqa: E402
from tests import workloads as wl  # noqa: E402


def cases():
    th = wl.emax_prior(200)
    yield "cfg1_emax_d", dict(model=1, mp=[1.0], theta=th, ub=10.0, K=3, crit=0, dt=None, m=1, cp=0, log_lo=None)
    th = wl.sigemax_prior(200)
    yield "cfg1b_sigemax_d", dict(model=2, mp=[1.0], theta=th, ub=1.0, K=8, crit=0, dt=None, m=1, cp=0, log_lo=None)
    th = wl.sigemax_prior(200, seed=2)
    yield "cfg2_sigemax_a_ed90", dict(model=2, mp=[1.0], theta=th, ub=1.0, K=6, crit=1, dt=wl.ded90_jacobians(th), m=1,
                                      cp=0, log_lo=None)
    th = wl.tpc_prior(200)
    yield "cfg3_tpc_d", dict(model=3, mp=[1.0], theta=th, ub=48.0, K=6, crit=0, dt=None, m=1, cp=0, log_lo=0.05)
    th = wl.tpc_prior(200, seed=4)
    yield "cfg4_vu_d", dict(model=4, mp=[0.1], theta=th, ub=24.0, K=3, crit=0, dt=None, m=3, cp=1, log_lo=None)


def main():
    out = {}
    for name, c in cases():
        pts, w = wl.jittered_swarm(12, c["K"], 0.0, c["ub"], seed=21, log_lo=c["log_lo"])
        omp = c["mp"] + ([c["m"]] if c["model"] == 4 else [])
        o = ko.Problem(c["model"], omp, c["theta"], crit=c["crit"], dt=c["dt"], m=c["m"], cp_id=c["cp"])
        out[name + "/points"], out[name + "/weights"] = pts, w
        out[name + "/objective"] = o.objective_batch(pts, w)
        dirs = np.linspace(0, c["ub"], 25)
        out[name + "/directions"] = dirs
        out[name + "/gateaux"] = o.gateauxderivative(pts[0], w[0], dirs)
        out[name + "/infomatrix_draw7"] = o.informationmatrix(pts[0], w[0], 7)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_fixtures.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
