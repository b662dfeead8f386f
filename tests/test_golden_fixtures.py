"""Committed golden vectors (tests/golden/).

reference_kats.json : known answers transcribed from the reference's tests/docs (provenance inside the file).
oracle_fixtures.npz : seeded inputs + CPU-oracle outputs written by tests/golden/make_fixtures.py.
CPU tests pin the oracle to both; the GPU tests check the CUDA engine against the same committed numbers,
so that parity does not silently depend on the oracle binary built on the day.
"""
import json
import os

import numpy as np
import pytest

import oracle as ko
from tests.golden.make_fixtures import cases

GOLD = os.path.join(os.path.dirname(__file__), "golden")
RTOL = 1e-10


@pytest.fixture(scope="module")
def kats():
    with open(os.path.join(GOLD, "reference_kats.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(GOLD, "oracle_fixtures.npz"))


def _oracle(c):
    omp = c["mp"] + ([c["m"]] if c["model"] == 4 else [])
    return ko.Problem(c["model"], omp, c["theta"], crit=c["crit"], dt=c["dt"], m=c["m"], cp_id=c["cp"])


def _gd(h, pts, w, dirs):
    assert not h.gateaux_prepare(pts, w)   # returns the singular flag
    return h.gateaux_eval(dirs)


def _engine(c):
    import kirstine_b200 as kb
    pw = np.full(len(c["theta"]), 1.0 / len(c["theta"]))
    return kb.ProblemHandle(c["model"], c["mp"], c["theta"], pw, c["crit"], c["dt"], (0.0,), (c["ub"],),
                            m=c["m"], cp_id=c["cp"])


# ---- CPU: oracle against the committed numbers ------------------------------------------------------------

def test_oracle_reference_kats(kats):
    k = kats["informationmatrix_emax"]
    o = ko.Problem(ko.EMAX, [k["sigma_squared"]], [k["theta"]])
    np.testing.assert_allclose(o.informationmatrix(k["design_points"], k["design_weights"]) * 16,
                               k["M_times_16"], rtol=1e-14)
    k = kats["tpc_locally_d_optimal"]
    o = ko.Problem(ko.TPC, [k["sigma"] ** 2], [k["theta"]])
    v = o.objective_batch([np.reshape(k["design_points"], (3, 1))], [k["design_weights"]])[0]
    assert v == pytest.approx(k["objective"], abs=k["atol"])
    k = kats["d_functional"]
    assert np.linalg.slogdet(np.array(k["matrix"]))[1] == pytest.approx(k["log_det"], rel=1e-15)


@pytest.mark.parametrize("name", [n for n, _ in cases()])
def test_oracle_reproduces_fixtures(fx, name):
    c = dict(cases())[name]
    o = _oracle(c)
    pts, w = fx[name + "/points"], fx[name + "/weights"]
    np.testing.assert_allclose(o.objective_batch(pts, w), fx[name + "/objective"], rtol=1e-13)
    np.testing.assert_allclose(o.gateauxderivative(pts[0], w[0], fx[name + "/directions"]), fx[name + "/gateaux"],
                               rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(o.informationmatrix(pts[0], w[0], 7), fx[name + "/infomatrix_draw7"], rtol=1e-13)


# ---- GPU: engine against the committed numbers ------------------------------------------------------------

@pytest.mark.gpu
def test_engine_reference_kats(kats):
    import kirstine_b200 as kb
    k = kats["informationmatrix_emax"]
    h = kb.ProblemHandle(1, [k["sigma_squared"]], [k["theta"]], [1.0], 0, None, (0.0,), (10.0,))
    np.testing.assert_allclose(h.informationmatrix(k["design_points"], k["design_weights"]) * 16,
                               k["M_times_16"], rtol=1e-14)
    k = kats["tpc_locally_d_optimal"]
    h = kb.ProblemHandle(3, [k["sigma"] ** 2], [k["theta"]], [1.0], 0, None, (0.0,), (48.0,))
    v = h.objective_batch([np.reshape(k["design_points"], (3, 1))], [k["design_weights"]])[0]
    assert v == pytest.approx(k["objective"], abs=k["atol"])
    # locally optimal sigmoid-Emax designs from the docs: s2 (end points and uniform weights fixed) is as good as
    # the free solution s1 to 1e-4, and the Gateaux derivative at s2 is <= ~0 everywhere (equivalence theorem)
    k = kats["sigemax_locally_optimal_designs"]
    h = kb.ProblemHandle(2, [k["sigma"] ** 2], [k["theta"]], [1.0], 0, None, (0.0,), (1.0,))
    f1 = h.objective_batch([np.reshape(k["s1_points"], (4, 1))], [k["s1_weights"]])[0]
    f2 = h.objective_batch([np.reshape(k["s2_points"], (4, 1))], [k["s2_weights"]])[0]
    assert abs(f2 - f1) < 1e-4
    gd = _gd(h, k["s2_points"], k["s2_weights"], np.linspace(0, 1, 1001))
    assert gd.max() < 1e-3
    k = kats["sigemax_discrete_prior_design"]
    h = kb.ProblemHandle(2, [k["sigma"] ** 2], k["thetas"], k["prior_weights"], 0, None, (0.0,), (1.0,))
    gd = _gd(h, k["points"], k["weights"], np.linspace(0, 1, 1001))
    assert gd.max() < 2e-2   # a 50-iteration PSO result: near-optimal (oracle: 1.4e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("name", [n for n, _ in cases()])
def test_engine_matches_fixtures(fx, name):
    c = dict(cases())[name]
    h = _engine(c)
    pts, w = fx[name + "/points"], fx[name + "/weights"]
    np.testing.assert_allclose(h.objective_batch(pts, w), fx[name + "/objective"], rtol=RTOL)
    gd = _gd(h, pts[0], w[0], fx[name + "/directions"])
    scale = np.abs(fx[name + "/gateaux"]).max()
    np.testing.assert_allclose(gd, fx[name + "/gateaux"], rtol=RTOL, atol=RTOL * scale)
    np.testing.assert_allclose(h.informationmatrix(pts[0], w[0], 7), fx[name + "/infomatrix_draw7"], rtol=RTOL)
