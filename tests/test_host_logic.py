"""Host-side mirror of the reference's data types (no GPU needed): validation and utilities."""
import numpy as np
import pytest

import kirstine_b200 as kb


def test_designmeasure_constructors():  # test/designmeasure.jl
    d = kb.DesignMeasure(([1], 0.2), ([42], 0.3), ([9], 0.5))
    assert kb.numpoints(d) == 3 and d.points.shape == (3, 1)
    assert d == kb.DesignMeasure([[1], [42], [9]], [0.2, 0.3, 0.5])
    with pytest.raises(ValueError, match="must be equal"):
        kb.DesignMeasure([[1], [2]], [1.0])
    with pytest.raises(ValueError, match="non-negative and sum to one"):
        kb.DesignMeasure([[1], [2]], [0.5, 0.6])
    with pytest.raises(ValueError, match="identical lengths"):
        kb.DesignMeasure([[1], [2, 3]], [0.5, 0.5])
    assert kb.one_point_design([42]) == kb.DesignMeasure([[42]], [1.0])
    assert np.allclose(kb.uniform_design([[1], [2], [3], [4]]).weights, 0.25)
    e = kb.equidistant_design(kb.DesignInterval(z=(0, 1)), 5)
    assert list(e.points.ravel()) == [0, 0.25, 0.5, 0.75, 1.0]
    with pytest.raises(ValueError, match="at least K = 2"):
        kb.equidistant_design(kb.DesignInterval(z=(0, 1)), 1)


def test_designinterval():  # test/designregion.jl
    dr = kb.DesignInterval(("dose", (0, 300)), ("time", (0, 20)))
    assert dr.N == 2 and dr.lowerbound == (0.0, 0.0) and dr.upperbound == (300.0, 20.0)
    assert kb.isinside([1, 2], dr) and not kb.isinside([301, 2], dr)
    with pytest.raises(ValueError, match="strictly larger"):
        kb.DesignInterval(a=(1, 1))


def test_sort_mixture_apportion():
    s = kb.sort_points(kb.uniform_design([[3, 4], [2, 1], [1, 1], [2, 3]]))
    assert s.points.tolist() == [[1, 1], [2, 1], [2, 3], [3, 4]]
    sw = kb.sort_weights(kb.DesignMeasure([[1], [2], [3]], [0.5, 0.2, 0.3]))
    assert sw.points.ravel().tolist() == [2, 3, 1]
    m = kb.mixture(0.25, kb.one_point_design([1]), kb.uniform_design([[2], [3]]))
    assert m.weights.tolist() == [0.25, 0.375, 0.375]
    with pytest.raises(ValueError, match="between 0 and 1"):
        kb.mixture(1.5, kb.one_point_design([1]), kb.one_point_design([2]))
    assert kb.apportion(np.array([0.5, 0.3, 0.2]), 10).tolist() == [5, 3, 2]
    assert kb.apportion(np.array([1 / 3, 1 / 3, 1 / 3]), 10).sum() == 10


def test_simplify():
    d = kb.DesignMeasure([[1], [2], [3]], [0.5, 0.001, 0.499])
    assert kb.numpoints(kb.simplify_drop(d, 0.01)) == 2
    dr = kb.DesignInterval(a=(0, 10))
    merged = kb.simplify_merge(kb.DesignMeasure([[1.0], [1.05], [5.0]], [0.25, 0.25, 0.5]), dr, 0.01)
    assert kb.numpoints(merged) == 2 and merged.points[0, 0] == pytest.approx(1.025)
    assert merged.weights[0] == pytest.approx(0.5)
    assert kb.simplify_merge(kb.one_point_design([1]), dr, 0.5) == kb.one_point_design([1])


def test_prior_sample_and_pso_validation():  # test/priorknowledge.jl, test/pso.jl:76-80
    with pytest.raises(ValueError, match="must be equal"):
        kb.PriorSample([kb.EmaxPar(1, 2, 3)], [0.5, 0.5])
    with pytest.raises(ValueError, match="sum to one"):
        kb.PriorSample([kb.EmaxPar(1, 2, 3), kb.EmaxPar(1, 2, 4)], [0.5, 0.6])
    pk = kb.PriorSample([kb.SigEmaxParameter(1, 2, 0.4, h) for h in range(1, 5)], [0.1, 0.3, 0.4, 0.2])
    assert kb.parameter_dimension(pk) == 4 and pk.parameter(2).h == 3
    with pytest.raises(ValueError, match="must be positive"):
        kb.Pso(iterations=-1, swarmsize=5, c1=3, c2=4)
    with pytest.raises(ValueError, match="at least 2"):
        kb.Pso(iterations=10, swarmsize=1, c1=3, c2=4)
    with pytest.raises(ValueError, match="c1 \\+ c2 > 4"):
        kb.Pso(iterations=10, swarmsize=5, c1=1, c2=1)
    with pytest.raises(ValueError, match="non-negative"):
        kb.Pso(iterations=10, swarmsize=5, c1=-1, c2=6)
    with pytest.raises(ValueError, match="steps must be >= 1"):
        kb.Exchange(optimizer_direction=kb.Pso(iterations=2, swarmsize=2), optimizer_weight=kb.Pso(iterations=2, swarmsize=2),
                    steps=0, candidate=kb.one_point_design([1]))


def test_trafo_constants_checks():  # test/transformation.jl
    pk = kb.PriorSample([kb.TPCParameter(4.298, 0.05884, 21.8)] * 3)
    assert kb.trafo_constants(kb.Identity(), pk) is None
    tc = kb.trafo_constants(kb.DeltaMethod(lambda p: [[p.s / p.a ** 2, -p.s / p.e ** 2, 1 / p.e - 1 / p.a]]), pk)
    assert tc.shape == (3, 1, 3)
    assert kb.codomain_dimension(kb.DeltaMethod(lambda p: np.eye(3)[:2]), pk) == 2
    with pytest.raises(ValueError, match="must have 3 columns"):
        kb.trafo_constants(kb.DeltaMethod(lambda p: [[1.0, 2.0]]), pk)
    with pytest.warns(UserWarning, match="more rows than cols"):
        kb.trafo_constants(kb.DeltaMethod(lambda p: np.ones((4, 3))), pk)


def test_unregistered_model_is_rejected():
    class MyModel:
        pass
    with pytest.raises(TypeError, match="not registered"):
        kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(a=(0, 1)), model=MyModel(),
                         covariate_parameterization=kb.CopyDose(), prior_knowledge=kb.PriorSample([kb.EmaxPar(1, 2, 3)]))


def test_shard_range_partitions_the_prior():
    for N, world in ((10, 3), (1_000_000, 8), (5, 8), (1000, 1)):
        ranges = [kb.shard_range(N, r, world) for r in range(world)]
        assert ranges[0][0] == 0 and ranges[-1][1] == N
        assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        assert all(0 <= j1 - j0 <= -(-N // world) for j0, j1 in ranges)


def test_apportion_pukelsheim_exhibit():  # test/designmeasure.jl:129-139 (Pukelsheim p. 310, Exhibit 12.2)
    w = np.array([1 / 6, 1 / 3, 1 / 2])
    ref = {7: [2, 2, 3], 8: [1, 3, 4], 9: [2, 3, 4], 10: [2, 3, 5], 11: [2, 4, 5], 12: [2, 4, 6]}
    for n, want in ref.items():
        assert kb.apportion(w, n).tolist() == want


def test_simplify_drop_reference_cases():  # test/designmeasure.jl:141-156
    d = kb.DesignMeasure([[1], [2], [3]], [1e-6, 0.5 - 5e-7, 0.5 - 5e-7])
    assert kb.simplify_drop(d, 1e-4) == kb.uniform_design([[2], [3]])
    assert kb.simplify_drop(d, 1e-6) == kb.uniform_design([[2], [3]])      # drop also on equality
    same = kb.simplify_drop(d, 1e-7)
    assert same == d and same is not d                                      # unchanged copy
    o = kb.one_point_design([42])
    o_simp = kb.simplify_drop(o, 1e-4)
    assert o_simp == o and o_simp is not o


def test_simplify_merge_reference_cases():  # test/designmeasure.jl:158-189
    d = kb.DesignMeasure([[1, 1], [5, 1], [3, 10]], [0.3, 0.1, 0.6])
    dr = kb.DesignInterval(a=(0, 100), b=(0, 100))
    ref = kb.DesignMeasure([[2, 1], [3, 10]], [0.4, 0.6])
    assert kb.simplify_merge(d, dr, 0.05) == ref
    assert kb.simplify_merge(d, dr, 0.04) == ref                           # merge also on equality
    same = kb.simplify_merge(d, dr, 0.03)
    assert same == d and same is not d
    o = kb.one_point_design([42, 42])
    o_simp = kb.simplify_merge(o, kb.DesignInterval(a=(0, 100), b=(1, 100)), 0.05)
    assert o_simp == o and o_simp is not o
    # unmerged points stay bit-identical (no round trip through the unit cube)
    dr1 = kb.DesignInterval(a=(0, 0.95))
    d1 = kb.uniform_design([[0.1], [0.5], [0.8], [0.81]])
    assert kb.simplify_merge(d1, dr1, 0) == d1
    merged = kb.simplify_merge(d1, dr1, 0.1)
    assert d1.points[:2].tolist() == merged.points[1:3].tolist()


def test_sort_points_reference_case():  # test/designmeasure.jl:192-200
    d = kb.DesignMeasure([[3, 4], [2, 1], [1, 1], [2, 3]], [0.4, 0.2, 0.3, 0.1])
    s = kb.sort_points(d)
    assert s == kb.DesignMeasure([[1, 1], [2, 1], [2, 3], [3, 4]], [0.3, 0.2, 0.1, 0.4])


def test_apply_transformation_values():  # test/design-common.jl:84-106
    M = np.array([[2.0, 1.0], [1.0, 3.0]])
    p = kb.EmaxPar(1, 2, 3)[:2]
    np.testing.assert_allclose(kb.apply_transformation(M, p, kb.Identity()), M)
    np.testing.assert_allclose(kb.apply_transformation(M, p, kb.DeltaMethod(lambda q: [[0, 1]])), [[2.5]], rtol=1e-15)
    np.testing.assert_allclose(kb.apply_transformation(M, p, kb.DeltaMethod(lambda q: [[2, 0], [-1, 1]])),
                               [[1.75, 2.0], [2.0, 3.0]], rtol=1e-14)
    # only the upper triangle is read
    junk = np.array([[2.0, 1.0], [99.0, 3.0]])
    np.testing.assert_allclose(kb.apply_transformation(junk, p, kb.DeltaMethod(lambda q: [[0, 1]])), [[2.5]], rtol=1e-15)


def test_implied_covariates():  # test/design-common.jl:205-216
    d = kb.DesignMeasure([[0], [1], [2]], [0.2, 0.3, 0.5])
    c = kb.implied_covariates(d, kb.EmaxModel(1), kb.CopyDose())
    assert [ci["dose"] for ci in c] == [0.0, 1.0, 2.0]
    cv = kb.implied_covariates(kb.DesignMeasure([[5], [20]], [0.25, 0.75]), kb.VUMod(0.1, 3), kb.EquiTime())
    assert cv[0]["time"] == [0.0, 5.0, 10.0] and cv[1]["time"] == [0.0, 20.0, 40.0]   # test/example-vector.jl:45-50
    cd = kb.implied_covariates(kb.DesignMeasure([[10, 0.5]], [1.0]), kb.DTRMod(1, 4), kb.LogEquidistantTimes(2.0))
    assert cd[0] == {"dose": 10.0, "time": [0.0, 0.5, 1.0, 2.0]}                       # docs/src/dtr.md:663-681
