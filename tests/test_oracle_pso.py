"""Pins the oracle's swarm-move geometry and PSO against the reference's unit tests."""
import numpy as np
import pytest

import oracle as ko

LB01, UB01 = [0.0], [1.0]


def test_move_designpoint_how_far():  # test/designregion.jl:67-84
    lb, ub = [1.0], [5.0]
    f = ko.move_designpoint_how_far
    assert f([3], [1.5], lb, ub) == 1
    assert f([3], [4.0], lb, ub) == 0.5
    assert f([3], [3.0], lb, ub) == 2 / 3
    assert f([3], [-1.5], lb, ub) == 1
    assert f([3], [-4.0], lb, ub) == 0.5
    assert f([3], [-3.0], lb, ub) == 2 / 3


def test_move_designpoint_clamps():  # test/designregion.jl:86-105 via move_add_v with K = 1
    lb, ub = [1.0], [5.0]
    for t, v, want in ((1, 1.5, 4.5), (0.5, 4.0, 5.0), (2 / 3, -3.0, 1.0)):
        pts, _ = ko.move_add_v([3.0], [1.0], t, [v], [0.0], lb, ub, [0])
        assert pts[0] == want


def test_move_handle_fixed():  # test/designmeasure-abstractpoint.jl:255-274
    va, vw = [0.25, 0.25, 0.25], [1 / 8, 3 / 8, 1 / 32]
    a, w = ko.move_handle_fixed(va, vw, [1, 0, 0], [0, 0, 0])
    assert list(w) == [0, 3 / 8, 1 / 32] and list(a) == va
    a, w = ko.move_handle_fixed(va, vw, [0, 0, 0], [1, 0, 0])
    assert list(w) == vw and list(a) == [0, 0.25, 0.25]
    a, w = ko.move_handle_fixed(va, vw, [0, 0, 1], [0, 0, 0])
    assert list(w) == [-1 / 8, 1 / 8, 1 / 32] and list(a) == va


def test_move_how_far():  # test/designmeasure-abstractpoint.jl:276-305
    pts, w = [0.5, 0.5, 0.5], [1 / 4, 1 / 4, 1 / 2]
    cases = [([0.1, 0.1, 0.1], [0.1, 0.1, 0.1], 1.0), ([1, 1, 0.1], [0.1, 0.1, 0.1], 0.5),
             ([1, -1, 0.1], [0.1, 0.1, 0.1], 0.5), ([0.1, 0.1, 0.1], [1, 1, 0.1], 0.25),
             ([0.1, 0.1, 0.1], [1, -1, 0.1], 0.25), ([1, 1, 0.1], [1, 1, 0.1], 0.25),
             ([1, -1, 0.1], [1, -1, 0.1], 0.25), ([1, 1, 0.1], [1, -1, 0.1], 0.25),
             ([1, -1, 0.1], [1, 1, 0.1], 0.25)]
    for va, vw, want in cases:
        assert ko.move_how_far(pts, w, va, vw, LB01, UB01) == want


def test_move_add_v():  # test/designmeasure-abstractpoint.jl:307-343
    lb, ub = [1.0], [3.0]
    pts, w = [1.5, 2, 2.5], [3 / 8, 4 / 8, 1 / 8]
    f = ko.move_add_v
    q = [0.25] * 3
    r = f(pts, w, 1.0, q, [1 / 32] * 3, lb, ub, [0, 0, 0])
    assert list(r[0]) == [1.75, 2.25, 2.75] and list(r[1]) == [13 / 32, 17 / 32, 1 / 16]
    r = f(pts, w, 1.0, [4, 0.25, 0.25], [1 / 32] * 3, lb, ub, [0, 0, 0])
    assert list(r[0]) == [3.0, 2.25, 2.75] and list(r[1]) == [13 / 32, 17 / 32, 1 / 16]
    r = f(pts, w, 1.0, [-2, 0.25, 0.25], [1 / 32] * 3, lb, ub, [0, 0, 0])
    assert list(r[0]) == [1.0, 2.25, 2.75]
    r = f(pts, w, 1.0, q, [1, -0.5, 0], lb, ub, [0, 0, 0])
    assert list(r[1]) == [1, 0, 0]
    r = f(pts, w, 1.0, q, [-1, 1 / 8, 2 / 8], lb, ub, [0, 0, 0])
    assert list(r[1]) == [0, 5 / 8, 3 / 8]
    r = f(pts, w, 1.0, q, [1 / 4, -1 / 4, 0], lb, ub, [0, 0, 1])
    assert list(r[1]) == [5 / 8, 2 / 8, 1 / 8]


def test_ap_add():  # test/designmeasure-abstractpoint.jl:229-253
    pts, w = [0.5, 0.5, 0.5], [1 / 4, 1 / 4, 1 / 2]
    nofix = [0, 0, 0]
    p1, w1, va1, vw1, t1 = ko.ap_add(pts, w, [0.1, 0.2, 0.3], [1 / 8, -1 / 8, 1 / 2], LB01, UB01, nofix, nofix)
    assert list(p1) == [0.6, 0.7, 0.8] and list(w1) == [3 / 8, 1 / 8, 1 / 2]
    assert list(va1) == [0.1, 0.2, 0.3] and list(vw1) == [1 / 8, -1 / 8, 1 / 2] and t1 == 1.0
    p2, w2, va2, vw2, t2 = ko.ap_add(pts, w, [-1, 1, 0.6], [1, 0, 1 / 2], LB01, UB01, nofix, nofix)
    assert list(p2) == [0, 1, 0.8] and list(w2) == [3 / 4, 1 / 4, 0]
    assert not np.any(va2) and not np.any(vw2)


def test_philox_known_answer():
    # Random123 kat_vectors: philox4x32-10, counter = key = 0 and all-ones
    assert ko.philox4x32(0, 0, 0, 0, 0) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert ko.philox4x32(0xffffffffffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff) == \
        [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    u = np.array([ko.uniform2(7, 1, i, 0, 0) for i in range(2000)]).ravel()
    assert 0 <= u.min() and u.max() < 1 and abs(u.mean() - 0.5) < 0.02


def _emax_problem():
    return ko.Problem(ko.EMAX, [1.0], [[1, 10, 5]])


def test_pso_emax_locally_optimal():  # test/solve-directmaximization.jl:13-70, example-emax.jl:49-61
    p = _emax_problem()
    pso = ko.Pso(p, S=50, K=3, lb=[0.0], ub=[10.0], seed=4711)
    pso.init([[0.0], [5.0], [10.0]], [1 / 3] * 3)
    st0 = pso.state()
    assert st0["n_eval"] == 50
    assert np.array_equal(st0["x_pts"][0].ravel(), [0, 5, 10])       # particle 1 = prototype
    assert np.all(st0["x_w"].sum(axis=1) == pytest.approx(1.0))
    assert np.all((st0["x_pts"] >= 0) & (st0["x_pts"] <= 10))
    assert not np.any(st0["v_pts"]) and not np.any(st0["v_w"])
    trace = [st0["fg"]]
    for _ in range(99):
        pso.tick()
        trace.append(pso.state()["fg"])
    st = pso.state()
    assert st["n_eval"] == 50 * 100                                    # Appendix B
    assert np.all(np.diff(trace) >= 0)                                 # monotone trace_fx
    order = np.argsort(st["g_pts"].ravel())
    np.testing.assert_allclose(st["g_pts"].ravel()[order], [0, 2.5, 10], atol=2e-2)
    np.testing.assert_allclose(st["g_w"][order], [1 / 3] * 3, atol=2e-2)
    sol = p.objective([0, 2.5, 10], [1 / 3] * 3)
    assert np.exp((st["fg"] - sol) / 3) > 0.9999                      # efficiency vs analytic optimum


def test_pso_fixed_components_stay_bit_identical():  # test/solve-directmaximization.jl:66-68,104-112
    p = _emax_problem()
    pso = ko.Pso(p, S=20, K=3, lb=[0.0], ub=[10.0], fixw=[1, 1, 1], fixp=[1, 0, 1], seed=3)
    pso.init([[0.0], [5.0], [10.0]], [1 / 3] * 3)
    for _ in range(15):
        pso.tick()
    st = pso.state()
    assert np.all(st["x_w"] == 1 / 3)
    assert np.all(st["x_pts"][:, 0, 0] == 0.0) and np.all(st["x_pts"][:, 2, 0] == 10.0)
    assert abs(st["g_pts"][1, 0] - 2.5) < 1e-2


def test_pso_error_cases():  # src/pso.jl:119-123, test/pso.jl:111-118
    th = [[1, 10, 5]]
    # a NaN prior weight makes every objective value NaN (0-weight/NaN rules of objective!)
    p = ko.Problem(ko.EMAX, [1.0], th, prior_weights=[float("nan")])
    pso = ko.Pso(p, S=5, K=3, lb=[0.0], ub=[10.0])
    with pytest.raises(RuntimeError, match="NaN on all"):
        pso.init([[0.0], [5.0], [10.0]], [1 / 3] * 3)
