"""Device-resident PSO and the solve()/Exchange host glue against the oracle PSO (same Philox stream)
and against the reference's known designs."""
import numpy as np
import pytest

import oracle as ko
from tests import workloads as wl

pytestmark = pytest.mark.gpu
kb = None


@pytest.fixture(scope="module", autouse=True)
def _engine():
    global kb
    import kirstine_b200
    kb = kirstine_b200
    assert kb.device_count() >= 1
    yield


def _emax_handles(theta=((1, 10, 5),)):
    theta = np.asarray(theta, float)
    pw = np.full(len(theta), 1.0 / len(theta))
    return kb.ProblemHandle(1, [1.0], theta, pw, 0, None, [0.0], [10.0]), ko.Problem(ko.EMAX, [1.0], theta)


def test_pso_init_matches_oracle_exactly():
    h, o = _emax_handles(wl.emax_prior(64))
    S, K = 37, 3
    proto = ([[0.0], [5.0], [10.0]], [1 / 3] * 3)
    assert h.pso_init(S, K, *proto, seed=4711) == 0
    po = ko.Pso(o, S, K, [0.0], [10.0], seed=4711)
    po.init(*proto)
    a, b = h.pso_state(), po.state()
    for key in ("x_pts", "x_w", "p_pts", "p_w", "v_pts", "v_w"):
        assert np.array_equal(a[key], b[key]), key      # same Philox stream, same arithmetic
    # random 3-point Emax designs are often ill-conditioned (cond >= 1e7); value parity at the 1e-10 bar is
    # asserted on conditioned designs in test_gpu_parity.py, here only the trajectory logic is under test
    np.testing.assert_allclose(a["fx"], b["fx"], rtol=1e-7)
    assert np.array_equal(a["g_pts"], b["g_pts"]) and a["n_eval"] == b["n_eval"] == S
    assert np.array_equal(a["x_pts"][0].ravel(), [0, 5, 10])


@pytest.mark.parametrize("fix", [None, ([1, 0, 0], [0, 0, 1]), ([0, 0, 1], [1, 0, 0]), ([1, 1, 1], [1, 0, 1])])
def test_pso_trajectory_matches_oracle(fix):
    """A few ticks: positions and velocities follow the oracle to rounding of fx-dependent choices."""
    h, o = _emax_handles(wl.emax_prior(200))
    S, K = 24, 3
    fixw, fixp = fix if fix else ([0, 0, 0], [0, 0, 0])
    proto = ([[0.0], [4.0], [10.0]], [0.3, 0.3, 0.4])
    assert h.pso_init(S, K, *proto, fixw=fixw, fixp=fixp, seed=99) == 0
    po = ko.Pso(o, S, K, [0.0], [10.0], fixw=fixw, fixp=fixp, seed=99)
    po.init(*proto)
    for it in range(6):
        tfx, tp, tw = h.pso_run(1)
        po.tick()
        a, b = h.pso_state(), po.state()
        for key in ("x_pts", "x_w", "v_pts", "v_w", "p_pts", "p_w", "g_pts", "g_w"):
            np.testing.assert_allclose(a[key], b[key], rtol=1e-12, atol=1e-14, err_msg="%s at tick %d" % (key, it))
        fin = np.isfinite(b["fx"])
        np.testing.assert_allclose(a["fx"][fin], b["fx"][fin], rtol=1e-7)
        assert tfx[0] == a["fg"] and np.array_equal(tp[0], a["g_pts"]) and np.array_equal(tw[0], a["g_w"])
        # fixed components stay bit-identical (test/solve-directmaximization.jl:66-68,104-112)
        for k in range(K):
            if fixw[k] or sum(fixw) >= K - 1:
                assert np.all(a["x_w"][:, k] == proto[1][k])
            if fixp[k]:
                assert np.all(a["x_pts"][:, k, 0] == proto[0][k][0])
    assert a["n_eval"] == S * 7


def test_pso_structural_invariants_and_trace():
    h, _ = _emax_handles(wl.emax_prior(500))
    S, K, iters = 50, 3, 100
    assert h.pso_init(S, K, [[0.0], [5.0], [10.0]], [1 / 3] * 3, seed=1) == 0
    fg0 = h.pso_state()["fg"]
    tfx, tp, tw = h.pso_run(iters - 1)
    st = h.pso_state()
    assert st["n_eval"] == S * iters
    assert np.all(np.diff(np.concatenate([[fg0], tfx])) >= 0)  # trace_fx non-decreasing
    assert tfx[-1] == st["fg"] == st["fp"].max()
    assert np.all((st["x_pts"] >= 0) & (st["x_pts"] <= 10))
    assert np.all(st["x_w"] >= 0) and np.allclose(st["x_w"].sum(axis=1), 1.0, atol=1e-12)
    assert np.all(tw.sum(axis=1) == pytest.approx(1.0))


def test_pso_error_statuses():  # src/pso.jl:119-123
    theta = np.array([[1.0, 10, 5]])
    h = kb.ProblemHandle(1, [1.0], theta, [float("nan")], 0, None, [0.0], [10.0])
    assert h.pso_init(5, 3, [[0.0], [5.0], [10.0]], [1 / 3] * 3) == 2      # NaN on all initial points


def test_solve_emax_locally_optimal_design():  # test/solve-directmaximization.jl:13-70, example-emax.jl:49-61
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 10)), model=kb.EmaxModel(1),
                          covariate_parameterization=kb.CopyDose(),
                          prior_knowledge=kb.PriorSample([kb.EmaxPar(e0=1, emax=10, ec50=5)]))
    kb.seed(4711)
    strategy = kb.DirectMaximization(optimizer=kb.Pso(iterations=50, swarmsize=50),
                                     prototype=kb.uniform_design([[0], [5], [10]]))
    d, r = kb.solve(dp, strategy, trace_state=True)
    res = kb.optimization_result(r)
    sol = kb.uniform_design([[0], [2.5], [10]])                             # analytic optimum (Dette et al. 2010)
    assert kb.efficiency(d, sol, dp) >= 0.9999
    assert np.all(np.diff(res.trace_fx) >= 0)
    assert len(res.trace_fx) == len(res.trace_x) == len(res.trace_state) == 50 and res.n_eval == 2500
    np.testing.assert_allclose(d.points.ravel(), [0, 2.5, 10], atol=5e-2)
    # fixed weights and points stay put (test/solve-directmaximization.jl:72-118)
    kb.seed(7)
    s2 = kb.DirectMaximization(optimizer=kb.Pso(iterations=20, swarmsize=20), prototype=kb.uniform_design([[0], [5], [10]]),
                               fixedweights=[0, 1, 2], fixedpoints=[0, 2])
    d2, r2 = kb.solve(dp, s2, trace_state=True)
    for st in kb.optimization_result(r2).trace_state:
        assert all(np.all(x.weights == 1 / 3) and x.points[0, 0] == 0 and x.points[2, 0] == 10 for x in st.x)
    assert abs(d2.points[1, 0] - 2.5) < 1e-2


def _sigemax_dp(prior, crit=None, trafo=None):
    return kb.DesignProblem(criterion=crit or kb.DCriterion(), region=kb.DesignInterval(dose=(0, 1)),
                            model=kb.SigEmaxModel(sigma=1), covariate_parameterization=kb.JustCopy("dose"),
                            prior_knowledge=prior, transformation=trafo)


def test_solve_reaches_reference_golden_designs():
    """>= 0.9999 efficiency against the designs the reference's docs pin (priors reproducible without Julia's RNG)."""
    # docs/src/locally-optimal.md:56-87
    dp = _sigemax_dp(kb.PriorSample([kb.SigEmaxParameter(e0=1, emax=2, ed50=0.4, h=5)]))
    gold = kb.DesignMeasure(([3.135435548592502e-5], 0.25000767275725305), ([0.3207541465365837], 0.25003991892306177),
                            ([0.48364781547861957], 0.24999249447251276), ([1.0], 0.24995991384717237))
    kb.seed(31415)
    s1, _ = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(iterations=50, swarmsize=50),
                                               prototype=kb.equidistant_design(kb.region(dp), 4)))
    assert kb.efficiency(s1, gold, dp) >= 0.9999
    # docs/src/locally-optimal.md:117-136: all weights and the end points fixed
    gold2 = kb.DesignMeasure(([0.0], 0.25), ([0.32036023546981274], 0.25), ([0.48370009596035385], 0.25), ([1.0], 0.25))
    s2, _ = kb.solve(dp, kb.DirectMaximization(optimizer=kb.Pso(iterations=20, swarmsize=50),
                                               prototype=kb.equidistant_design(kb.region(dp), 4),
                                               fixedweights=[0, 1, 2, 3], fixedpoints=[0, 3]))
    assert kb.efficiency(s2, gold2, dp) >= 0.9999
    assert np.all(s2.weights == 0.25) and s2.points[0, 0] == 0 and s2.points[3, 0] == 1
    # docs/src/discrete-prior.md:59-95: four-point prior with weights
    prior = kb.PriorSample([kb.SigEmaxParameter(e0=1, emax=2, ed50=0.4, h=h) for h in range(1, 5)], [0.1, 0.3, 0.4, 0.2])
    dpd = _sigemax_dp(prior)
    gold3 = kb.DesignMeasure(([0.0], 0.17962034176078864), ([0.04461948295716912], 0.0934736269518622),
                             ([0.25627704567921966], 0.23934177101306972), ([0.4958739526513957], 0.23846947150228448),
                             ([1.0], 0.24909478877199498))
    s3, _ = kb.solve(dpd, kb.DirectMaximization(optimizer=kb.Pso(iterations=50, swarmsize=100),
                                                prototype=kb.equidistant_design(kb.region(dpd), 8)),
                     maxweight=1e-3, maxdist=1e-2)
    assert kb.efficiency(s3, gold3, dpd) >= 0.9999
    # equivalence theorem at the solution
    gd = kb.gateauxderivative(s3, [kb.one_point_design([x]) for x in np.linspace(0, 1, 101)], dpd)
    assert gd.max() < 5e-3


def test_public_evaluators_semantics():  # test/designproblem.jl:75-167
    dp = _sigemax_dp(kb.PriorSample([kb.SigEmaxParameter(1, 2, 0.4, 5)]))
    good = kb.equidistant_design(kb.region(dp), 5)
    sing = kb.one_point_design([0.5])
    assert kb.objective(sing, dp) == -np.inf
    assert kb.efficiency(sing, good, dp) == 0.0 and kb.efficiency(good, sing, dp) == np.inf
    assert np.isnan(kb.efficiency(sing, sing, dp))
    assert kb.efficiency(good, good, dp) == 1.0
    dpa = _sigemax_dp(dp.pk, crit=kb.ACriterion())
    assert kb.efficiency(good, kb.equidistant_design(kb.region(dp), 6), dpa) == \
        kb.efficiency(good, kb.equidistant_design(kb.region(dp), 6), dp)       # criterion is ignored
    dp1 = _sigemax_dp(dp.pk, trafo=kb.DeltaMethod(lambda p: [[0, 0, 1, 0]]))
    with pytest.raises(ValueError, match="dimensions of transformed parameters"):
        kb.efficiency(good, good, dp, dp1)
    n = 10
    want = 2 * (np.log(n) - 1 + np.log(2 * np.pi)) + 0.5 * kb.objective(good, dp)
    assert kb.shannon_information(good, dp, n) == pytest.approx(want, rel=1e-14)
    with pytest.raises(RuntimeError, match="one-point directions"):
        kb.gateauxderivative(good, [good], dp)
    assert np.all(np.isnan(kb.gateauxderivative(sing, [kb.one_point_design([0.1])], dp)))
    M = kb.informationmatrix(kb.DesignMeasure(([0], 0.25), ([5], 0.75)), kb.EmaxModel(1), kb.CopyDose(),
                             kb.EmaxPar(1, 10, 5), kb.FisherMatrix())
    np.testing.assert_allclose(M, np.array([[16, 6, -6], [6, 3, -3], [-6, -3, 3]]) / 16, rtol=1e-14)


def test_gateauxderivative_grid_export(tmp_path):
    """Headless form of plot_gateauxderivative (src/plot.jl:103-187; sizes as checked in test/plot.jl:17-60)."""
    import kirstine_b200 as kb
    # 1-d: Emax, locally optimal design on [0, 1] (test/example-emax.jl:49-61)
    e0, emax, ec50 = 0.0, 1.0, 0.2
    a, b = 0.0, 1.0
    xstar = (a * (b + ec50) + b * (a + ec50)) / (a + b + 2 * ec50)
    d = kb.uniform_design([[a], [xstar], [b]])
    dp = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 1)),
                          prior_knowledge=kb.PriorSample([kb.EmaxPar(e0, emax, ec50)]), model=kb.EmaxModel(1),
                          covariate_parameterization=kb.CopyDose())
    x, gd = kb.gateauxderivative_grid(d, dp)
    assert len(x) == len(gd) == 101 + 3 and np.all(np.diff(x) >= 0)
    assert gd.max() < 1e-9                                    # equivalence theorem at the optimal design
    csv = tmp_path / "gd.csv"
    x2, gd2 = kb.gateauxderivative_grid(d, dp, n_grid=11, csv_path=str(csv))
    assert len(x2) == 14                                      # grid with 11 points plus the three design points
    tab = np.loadtxt(csv, delimiter=",", skiprows=1)
    assert tab.shape == (14, 2) and open(csv).readline().strip() == "dose,gateauxderivative"
    np.testing.assert_array_equal(tab[:, 1], gd2)
    with pytest.raises(ValueError, match="single integer"):
        kb.gateauxderivative_grid(d, dp, n_grid=(3, 3))
    with pytest.raises(ValueError, match=">= 2"):
        kb.gateauxderivative_grid(d, dp, n_grid=1)
    # 2-d: dose-time-response with CopyBoth (test/example-dtr.jl)
    th = wl.dtr_prior(50)
    dp2 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 100), time=(0, 24)),
                           prior_knowledge=kb.PriorSample(th), model=kb.DTRMod(1, 1),
                           covariate_parameterization=kb.CopyBoth())
    d2 = kb.uniform_design([[100, 1], [100, 4], [50, 8], [20, 12], [100, 24], [60, 2]])
    rx, ry, g2 = kb.gateauxderivative_grid(d2, dp2)
    assert rx.shape == (51,) and ry.shape == (51,) and g2.shape == (51, 51) and np.all(np.isfinite(g2))
    one = kb.gateauxderivative(d2, [kb.one_point_design([rx[7], ry[30]])], dp2)[0]
    assert g2[7, 30] == pytest.approx(one, rel=1e-12)
    rx, ry, g3 = kb.gateauxderivative_grid(d2, dp2, n_grid=(4, 6))
    assert g3.shape == (4, 6)
    with pytest.raises(ValueError, match="2-tuple"):
        kb.gateauxderivative_grid(d2, dp2, n_grid=5)


def test_pso_long_run_in_pieces_equals_one_run():
    """cfg-5-shaped swarm (1024 particles, K = 8) for 500 iterations, the iteration count BASELINE.json quotes:
    monotone trace, and running it as 5 x 100 iterations gives the same bits as one run of 500 (the state lives
    on the device between calls, the random stream is keyed on the iteration counter)."""
    import kirstine_b200 as kb
    th = wl.sigemax_prior(2000, seed=5)
    pw = np.full(len(th), 1.0 / len(th))
    proto = (np.linspace(0, 1, 8)[:, None], np.full(8, 1 / 8))
    h1 = kb.ProblemHandle(2, [1.0], th, pw, 0, None, (0.0,), (1.0,))
    assert h1.pso_init(1024, 8, *proto, seed=7) == 0
    t1 = h1.pso_run(500)[0]
    h5 = kb.ProblemHandle(2, [1.0], th, pw, 0, None, (0.0,), (1.0,))
    assert h5.pso_init(1024, 8, *proto, seed=7) == 0
    t5 = np.concatenate([h5.pso_run(100)[0] for _ in range(5)])
    assert np.array_equal(t1, t5)
    assert np.all(np.diff(t1) >= 0) and np.isfinite(t1[-1])
    s1, s5 = h1.pso_state(), h5.pso_state()
    for key in ("x_pts", "x_w", "v_pts", "v_w", "p_pts", "p_w", "g_pts", "g_w", "fx", "fp"):
        assert np.array_equal(s1[key], s5[key]), key
    assert s1["n_eval"] == 1024 * 501
    # the solved design is (near) optimal: Gateaux derivative <= ~0 over the region
    g = kb.DesignMeasure(s1["g_pts"], s1["g_w"] / s1["g_w"].sum())
    assert not h1.gateaux_prepare(g.points, g.weights)
    assert h1.gateaux_eval(np.linspace(0, 1, 501)).max() < 1e-2


@pytest.mark.parametrize("family", ["tpc", "vu", "dtr"])
def test_direction_pso_trajectory_matches_oracle(family):
    """The Exchange strategy's inner swarm (src/solve-exchange.jl:133-149): one-point directions, objective = Gateaux
    derivative at the prepared candidate.  Engine state against the oracle's on the shared random stream."""
    import kirstine_b200 as kb
    if family == "tpc":
        model, mp, omp, th, lb, ub, kw, okw = 3, [1.0], [1.0], wl.tpc_prior(150, seed=8), [0.0], [48.0], {}, {}
        cand = ([[0.3], [2.0], [20.0]], [0.3, 0.3, 0.4])
    elif family == "vu":
        model, mp, omp, th, lb, ub = 4, [0.1], [0.1, 3], wl.tpc_prior(150, seed=9), [0.0], [24.0]
        kw, okw = dict(m=3, cp_id=1), dict(m=3, cp_id=ko.CP_EQUITIME)
        cand = ([[2.0], [8.0], [20.0]], [0.3, 0.3, 0.4])
    else:
        model, mp, omp, th, lb, ub = 5, [1.0], [1.0], wl.dtr_prior(100, seed=3), [0.0, 0.0], [100.0, 24.0]
        kw, okw = dict(m=1, cp_id=2), dict(m=1, d=2, cp_id=2)
        pts, w = wl.jittered_swarm_nd(1, 7, lb, ub, seed=4)
        cand = (pts[0], w[0])
    pw = np.full(len(th), 1.0 / len(th))
    h = kb.ProblemHandle(model, mp, th, pw, 0, None, lb, ub, **kw)
    o = ko.Problem(model, omp, th, pw, **okw)
    assert not h.gateaux_prepare(*cand)
    gconst = o.gateaux_constants(*cand)
    S = 25
    protos = np.asarray(cand[0], float).reshape(-1, 1, len(lb))[:3]
    pwts = np.ones((len(protos), 1))
    assert h.pso_init(S, 1, protos, pwts, fixw=[0], fixp=[0], seed=5, mode=kb._ffi.PSO_MODE_GATEAUX) == 0
    po = ko.Pso(o, S, 1, lb, ub, seed=5, mode=1, gconst=gconst)
    po.init(protos, pwts)
    for it in range(8):
        a, b = h.pso_state(), po.state()
        for key in ("x_pts", "v_pts", "p_pts", "g_pts"):
            np.testing.assert_allclose(a[key], b[key], rtol=1e-11, atol=1e-13, err_msg="%s at tick %d" % (key, it))
        np.testing.assert_allclose(a["fx"], b["fx"], rtol=1e-8, atol=1e-9 * np.abs(b["fx"]).max())
        assert np.all(a["x_w"] == 1.0)
        h.pso_run(1)
        po.tick()


def test_profile_stages_time_one_launch_each(monkeypatch):
    """kst_profile_enable(h, stage): events around the evaluation kernel (1), the move (2) or the best update (3) of the
    three-launch PSO step -- one pair per iteration each, none for the other stages; a stage outside 0..3 is refused."""
    monkeypatch.setenv("KST_NO_FUSED", "1")   # the persistent kernel has no separate launches to time
    h, _ = _emax_handles(wl.emax_prior(2000))
    assert h.pso_init(64, 3, [[0.0], [5.0], [10.0]], [1 / 3] * 3, seed=3) == 0
    for stage in (1, 2, 3):
        h.profile_enable(stage)
        h.pso_run(4)
        ms, n = h.profile_read()
        assert n == 4 and 0.0 < ms < 50.0, (stage, ms, n)
    h.profile_enable(0)
    h.pso_run(2)
    assert h.profile_read()[1] == 0
    with pytest.raises(kb.EngineError, match="stage"):
        h.profile_enable(4)


@pytest.mark.parametrize("S", [300, 1000])
def test_pso_many_particles_cross_cta_best_matches_oracle(S, monkeypatch):
    """Swarms of more than 128 particles: the best update runs as several CTAs whose candidates the last-arriving one
    merges (k_pso_best), and the move takes its particles in groups.  Every prototype appears twice, 150 indices
    apart, so equal objective values meet across CTAs and the lower index has to win as in the reference's sequential
    pso_update_best! (src/pso.jl:181-193); positions, velocities and bests follow the oracle tick by tick."""
    monkeypatch.setenv("KST_NO_FUSED", "1")
    h, o = _emax_handles(wl.emax_prior(300, seed=11))
    K, half = 3, 150
    rng = np.random.default_rng(7)
    pts = np.sort(rng.uniform(0.0, 10.0, (half, K, 1)), axis=1)
    w = rng.dirichlet(np.ones(K), half)
    pts, w = np.concatenate([pts, pts]), np.concatenate([w, w])      # particle i + 150 == particle i
    assert h.pso_init(S, K, pts, w, seed=5) == 0
    po = ko.Pso(o, S, K, [0.0], [10.0], seed=5)
    po.init(pts, w)
    a, b = h.pso_state(), po.state()
    for key in ("x_pts", "x_w", "p_pts", "p_w", "g_pts", "g_w"):
        # (the random particles' weights go through log: device and host libm may differ in the last bit)
        np.testing.assert_allclose(a[key], b[key], rtol=1e-14, atol=0, err_msg=key)
    assert np.array_equal(a["fx"][:half], a["fx"][half:2 * half])     # the duplicates really tie
    for it in range(4):
        h.pso_run(1)
        po.tick()
        a, b = h.pso_state(), po.state()
        for key in ("x_pts", "x_w", "v_pts", "v_w", "p_pts", "p_w", "g_pts", "g_w"):
            np.testing.assert_allclose(a[key], b[key], rtol=1e-12, atol=1e-14, err_msg="%s at tick %d" % (key, it))
        assert a["fg"] == a["fp"].max()
    assert a["n_eval"] == S * 5
