"""The acceptance rules of tests/referee.py:assert_parity, on synthetic numbers (no GPU): what the GPU parity
tests can and cannot let through."""
import numpy as np
import pytest

from tests import referee as rf

INF = np.inf


@pytest.mark.parametrize("g,o", [(1.0, -INF), (-INF, 1.0), (np.nan, 1.0), (1.0, np.nan), (INF, -INF), (np.nan, -INF)])
def test_inf_nan_classes_must_match(g, o):
    with pytest.raises(AssertionError):
        rf.assert_parity([g], [o], lambda i: 1.0)
    with pytest.raises(AssertionError):     # a merely ill-conditioned design does not excuse it either
        rf.assert_parity([g], [o], lambda i: 1.0, cond_fn=lambda i: 1e9)
    rf.assert_parity([g], [o], lambda i: 1.0, cond_fn=lambda i: 1e13)   # numerically singular: rounding luck
    assert rf.last_report["singular_luck"] == 1


def test_matching_classes_pass():
    rf.assert_parity([-INF, np.nan, INF], [-INF, np.nan, INF], lambda i: 1.0)
    assert rf.last_report["class_match"] == 3


def test_well_conditioned_designs_get_no_slack():
    ok = dict(cond_fn=lambda i: 50.0)
    rf.assert_parity([1.0 + 5e-11], [1.0], lambda i: 1.0, **ok)                  # within the bar of the oracle
    rf.assert_parity([1.0 + 1e-9], [1.0], lambda i: 1.0 + 1e-9, **ok)           # oracle off, engine on the referee
    assert rf.last_report["referee"] == 1
    with pytest.raises(AssertionError):                                          # both off: no slack below cond 1e6
        rf.assert_parity([1.0 + 1e-9], [1.0], lambda i: 1.0 + 5e-9, **ok)


def test_ill_conditioned_designs_get_a_slack_of_eight():
    ill = dict(cond_fn=lambda i: 1e8)
    rf.assert_parity([1.0 + 1e-9], [1.0], lambda i: 1.0 + 5e-9, **ill)          # engine error 4e-9 <= 8 x 5e-9
    assert rf.last_report["slack"] == 1
    with pytest.raises(AssertionError):
        rf.assert_parity([1.0 + 1e-7], [1.0], lambda i: 1.0 + 5e-9, **ill)      # 9.5e-8 > 8 x 5e-9
    rf.assert_parity([1.0 + 1e-9], [1.0], lambda i: 1.0, **ill)                  # oracle exact by luck: inside the
    assert rf.last_report["cond_bound"] == 1                                     # cond * 2^-53 = 1.1e-8 bound
    with pytest.raises(AssertionError):                                          # beyond it: no excuse
        rf.assert_parity([1.0 + 5e-8], [1.0], lambda i: 1.0, **ill)


def test_scaled_condition_number():
    # Emax, locally: three well-separated doses are well conditioned once the columns are scaled
    th = np.array([[1.0, 10.0, 5.0]])
    good = rf.cond_max(1, [1.0], [[0.0], [2.5], [10.0]], [1 / 3] * 3, th)
    near = rf.cond_max(1, [1.0], [[0.0], [2.5], [2.5 + 1e-7]], [1 / 3] * 3, th)
    assert good < 1e4 and near > 1e12
    assert rf.cond_max(1, [1.0], [[2.5]], [1.0], th) > 1e12 or np.isinf(rf.cond_max(1, [1.0], [[2.5]], [1.0], th))


def test_cond_bound_rule_and_transformed_matrix():
    ill = dict(cond_fn=lambda i: 1e9)       # cond * 2^-53 = 1.1e-7 relative
    rf.assert_parity([1.0 + 5e-9], [1.0], lambda i: 1.0 + 2e-10, **ill)        # oracle lucky (2e-10), engine 4.8e-9
    assert rf.last_report["cond_bound"] == 1
    with pytest.raises(AssertionError):
        rf.assert_parity([1.0 + 5e-7], [1.0], lambda i: 1.0 + 2e-10, **ill)    # beyond the forward error bound
    # a dense DeltaMethod Jacobian on a compartment-model design: M is harmless once scaled, DT M^-1 DT' is close to
    # rank one because the columns of M^-1 differ by many orders of magnitude
    th = np.array([[4.3, 0.06, 21.8]])
    pts, w = [[3.0], [4.4], [13.7]], [0.3, 0.3, 0.4]
    q = np.linalg.qr(np.random.default_rng(0).normal(size=(3, 3)))[0]
    plain = rf.cond_max(4, [0.1], pts, w, th, m=3)
    with_dt = rf.cond_max(4, [0.1], pts, w, th, m=3, dt=q[None, :2, :])
    assert plain < 1e3 and with_dt > 1e4 * plain
