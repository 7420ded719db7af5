"""Pins the oracle's internal building blocks (the ones with no device-side twin) against
the reference's golden values: Sharder (sharder_test.cc:86-190), the linear-time and
diagonal trust-region solvers (trust_region_test.cc:53-500,963-1066), one-iteration Ruiz/L2
rescaling (sharded_optimization_utils_test.cc:303-357), the Lagrangian gradients
(:402-510) and the singular-value estimate (:517-592)."""
import math

import numpy as np
import pytest
import scipy.sparse as sp

import fixtures
import oracle_binding as ob
import pdlp_b200

INF = np.inf


# ---- Sharder ----
def verify_sharder(starts, target, masses):
    starts = list(starts)
    n = len(masses)
    assert starts[0] == 0 and starts[-1] == n
    num_shards = len(starts) - 1
    assert 1 <= num_shards <= 2 * target
    shard_mass = [sum(masses[starts[s]:starts[s + 1]]) for s in range(num_shards)]
    for s in range(num_shards):
        assert starts[s + 1] - starts[s] >= 1 and shard_mass[s] >= 1
    total, mx = sum(masses), max(masses)
    upper = max(mx, -(-mx // 2) + -(-total // target))
    lower = total // target - -(-mx // 2)
    for s in range(num_shards):
        assert shard_mass[s] <= upper
        if s + 1 < num_shards:
            assert shard_mass[s] >= lower


def test_sharder_from_matrix():
    # column masses of TestSparseMatrix are nnz+1 = [4, 2, 2, 3] (sharder.cc:72-76)
    verify_sharder(ob.shard_starts(4, 2, [4, 2, 2, 3]), 2, [4, 2, 2, 3])


@pytest.mark.parametrize("n,shards,want", [(10, 3, None), (5, 7, [0, 1, 2, 3, 4, 5]), (5, 1_000_000_000, [0, 1, 2, 3, 4, 5]), (5, 1, [0, 5]), (1, 5, [0, 1])])
def test_uniform_sharder(n, shards, want):
    starts = ob.shard_starts(n, shards)
    if want is not None:
        assert list(starts) == want
    verify_sharder(starts, min(shards, 7), [1] * n)


def test_uniform_sharder_zero_elements():
    assert list(ob.shard_starts(0, 3)) == [0]


# ---- trust region ----
TR_CASES = [
    # name, lb, ub, center, objective, weights, radius, solution, objective value, tol on objective for the diagonal solver
    ("no_bounds", [-INF, -INF], [INF, INF], [2.0, -5.0], [1.0, 1.0], [1, 1], math.sqrt(2.0), [1.0, -6.0], -2.0, 1e-6),
    ("with_bounds", [2.0, -INF, -INF], [INF, INF, INF], [2.0, -5.0, 1.0], [1.0, -1.0, 1.0], [1, 1, 1], math.sqrt(2.0), [2.0, -4.0, 0.0], -2.0, 1e-6),
    ("at_bounds", [2.0, -INF], [INF, -5.0], [2.0, -5.0], [1.0, -1.0], [1, 1], 1.0, [2.0, -5.0], 0.0, 1e-6),
    ("inactive_radius", [2.0, -INF, 0.5], [INF, -5.0, INF], [2.0, -5.0, 1.0], [1.0, -1.0, 1.0], [1, 1, 1], 1.0, [2.0, -5.0, 0.5], -0.5, 1e-6),
    ("zero_radius", [2.0, -INF, 0.5], [INF, -5.0, INF], [2.0, -5.0, 1.0], [1.0, -1.0, 1.0], [1, 1, 1], 0.0, [2.0, -5.0, 1.0], 0.0, 1e-6),
    ("infinite_radius", [2.0, -INF, 0.5], [INF, -5.0, INF], [2.0, -5.0, 1.0], [1.0, -1.0, 1.0], [1, 1, 1], INF, [2.0, -5.0, 0.5], -0.5, 1e-6),
    ("mixed_objective", [1.0, 0.0], [INF, INF], [2.0, 1.0], [2.0, 1.0], [1, 1], math.sqrt(1.25), [1.0, 0.5], -2.5, 2e-6),
    ("zero_objective", [-INF], [INF], [2.0], [0.0], [1], 1.0, [2.0], 0.0, 1e-6),
    ("weights_no_bounds", [-INF, -INF], [INF, INF], [2.0, -5.0], [1.0, 2.0], [1.0, 2.0], math.sqrt(3.0), [1.0, -6.0], -3.0, 1e-5),
    ("weights_with_bounds", [2.0, -INF, -INF], [INF, INF, INF], [2.0, -5.0, 1.0], [0.5, -2.0, 3.0], [0.5, 2.0, 3.0], math.sqrt(5.0), [2.0, -4.0, 0.0], -5.0, 1e-5),
    ("weights_hits_bound", [1.0, 0.0], [INF, INF], [2.0, 1.0], [1.0, 2.0], [0.5, 2.0], 1.0, [1.0, 0.5], -2.0, 1e-6),
    ("weights_large", [1.0, 0.0], [INF, INF], [2.0, 1.0], [1000.0, 2.0], [500.0, 2.0], math.sqrt(500.5), [1.0, 0.5], -1001.0, 1e-6),
]


@pytest.mark.parametrize("case", TR_CASES, ids=[c[0] for c in TR_CASES])
@pytest.mark.parametrize("shards", [1, 2])
def test_linear_time_trust_region(case, shards):
    _, lb, ub, center, obj, w, radius, want, want_obj, _ = case
    sol, _, value = ob.trust_region(obj, lb, ub, center, w, radius, num_shards=shards)
    np.testing.assert_allclose(sol, want, rtol=0, atol=1e-13)
    assert abs(value - want_obj) <= 4 * np.spacing(max(abs(want_obj), 1.0))


@pytest.mark.parametrize("case", TR_CASES, ids=[c[0] for c in TR_CASES])
def test_diagonal_trust_region_on_lps(case):
    _, lb, ub, center, obj, w, radius, want, want_obj, obj_tol = case
    sol, _, value = ob.trust_region(obj, lb, ub, center, w, radius, diag=np.zeros(len(obj)), num_shards=2, tol=1e-8 if case[0] == "no_bounds" else 1e-6)
    np.testing.assert_allclose(sol, want, rtol=0, atol=1e-6)
    assert abs(value - want_obj) <= obj_tol


def one_dim_qp_data(primal_weight):
    # GenerateTestQpProblemData (trust_region_test.cc:859-878)
    return dict(obj=[2.0, -1.0], diag=[2.0, 0.0], center=[0.0, -1.0], lb=[-INF, -INF], ub=[INF, INF], w=[0.5 * primal_weight, 0.5 / primal_weight])


def test_diagonal_trust_region_one_dim_qp():
    d = one_dim_qp_data(1.0)
    sol, step, value = ob.trust_region(d["obj"], d["lb"], d["ub"], d["center"], d["w"], 0.5, diag=d["diag"], num_shards=2, tol=1e-6)
    np.testing.assert_allclose(sol, [-0.5, -0.5], atol=1e-6)
    assert abs(step - 4.0) <= 4e-6 and abs(value + 1.25) <= 1e-6
    d = one_dim_qp_data(100.0)
    _, step, _ = ob.trust_region(d["obj"], d["lb"], d["ub"], d["center"], d["w"], math.sqrt(2705.0 / 2) * (5.0 / 13), diag=d["diag"], num_shards=2, tol=1e-6)
    assert abs(step - 1.0) <= 1e-6
    d = one_dim_qp_data(0.01)
    sol, step, value = ob.trust_region(d["obj"], d["lb"], d["ub"], d["center"], d["w"], 0.71063, diag=d["diag"], num_shards=2, tol=1e-6)
    np.testing.assert_allclose(sol, [-0.99950025, -0.9], atol=1e-6)
    assert abs(step - 0.2) <= 1e-6 and abs(value + 1.0999996) <= 1e-6


# ---- one-iteration rescaling goldens ----
def test_linf_ruiz_one_iteration():
    p = ob.Problem(fixtures.make_test_lp(), 2, 2)
    rs, cs = p.scaling_vectors(0, 1, [1, 2, 1, 3.0], [0, 1, 2, -1.0])
    np.testing.assert_array_equal(rs, [1 / math.sqrt(2), 1.0, 1.0, 1.0])
    np.testing.assert_array_equal(cs, [0.0, 1.0, 2.0 / 3.0, -1.0 / math.sqrt(3.0)])


def test_l2_rescaling_one_iteration():
    p = ob.Problem(fixtures.make_test_lp(), 2, 2)
    rs, cs = p.scaling_vectors(1, 1, [1, 2, 1, 3.0], [0, 1, 2, -1.0])
    np.testing.assert_allclose(rs, [1.0 / 3.0 ** 0.5, 1.0, 1.0, 3.0 / 90.0 ** 0.25], rtol=4e-16)
    np.testing.assert_allclose(cs, [0.0, 1.0, 2.0 / 101 ** 0.25, -1.0 / 13.0 ** 0.25], rtol=4e-16)


def test_l2_rescaling_non_square():
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(2, 1)
    qp.constraint_matrix = sp.csc_matrix(np.array([[2.0, 3.0]]))
    rs, cs = ob.Problem(qp, 2, 2).scaling_vectors(1, 1, [1.0], [1.0, 1.0])
    np.testing.assert_allclose(rs, [1.0 / 13.0 ** 0.25], rtol=4e-16)
    np.testing.assert_allclose(cs, [1.0 / math.sqrt(2.0), 1.0 / math.sqrt(3.0)], rtol=4e-16)


# ---- Lagrangian gradients ----
def test_lagrangian_gradients_lp():
    p = ob.Problem(fixtures.make_test_lp(), 2, 2)
    pg, dg, vals = p.lagrangian_gradients([0.0, 0.0, 0.0, 3.0], [-1.0, 0.0, 1.0, 1.0])
    np.testing.assert_array_equal(pg, [5.5 - 2.0, -2.0 + 1.0, -1.0 - 0.5, 1.0 + 3.0])
    assert vals[0] == 3.0 + 9.0
    np.testing.assert_array_equal(dg, [12.0 - 6.0, 7.0, -4.0, -1.0 + 3.0])
    assert vals[1] == 12.0 * -1.0 + -4.0 * 1.0 + -1.0 * 1.0


def test_dual_gradient_two_sided_constraints():
    qp = fixtures.make_test_lp()
    qp.constraint_lower_bounds[0] = 4
    qp.constraint_lower_bounds[1] = 5
    qp.constraint_upper_bounds[2] = -1
    _, dg, vals = ob.Problem(qp, 2, 2).lagrangian_gradients([0.0, 0.0, 0.0, 3.0], [0.0, 0.0, 0.0, -1.0])
    np.testing.assert_array_equal(dg, [0.0, 5.0, -1.0, 4.0])
    assert vals[1] == -1.0


def test_lagrangian_gradients_qp():
    p = ob.Problem(fixtures.diagonal_qp1(), 2, 2)
    pg, dg, vals = p.lagrangian_gradients([1.0, 2.0], [-2.0])
    np.testing.assert_array_equal(pg, [5.0, 3.0])
    assert vals[0] == 4.0 - 3.0 + 2.0 * 3.0
    np.testing.assert_array_equal(dg, [1.0 - 3.0])
    assert vals[1] == -2.0


# ---- power method ----
def test_sigma_max_test_lp():
    # the exact value; the reference's mt19937/absl::Gaussian start vector is not
    # reproducible outside abseil ("parity unpinned"), the estimate's accuracy is.
    v, iters = ob.Problem(fixtures.make_test_lp(), 2, 2).sigma_max(0.01, 0.001, 1)
    assert abs(v - 4.76945) < 0.01 and iters < 300


def test_sigma_max_diagonal():
    qp = fixtures.make_test_lp()
    qp.constraint_matrix = sp.csc_matrix(np.diag([2.0, 1.0, -3.0, -1.0]))
    v, iters = ob.Problem(qp, 2, 2).sigma_max(0.01, 0.001, 1)
    assert abs(v - 3.0) < 1e-4 and iters < 300


@pytest.mark.parametrize("x, y, want", [
    # x_1 at its bound (-2): the matrix loses column 1 (:529-541); y_1 = 0 on a one-sided row: it loses row 1 (:544-557).
    ([0.0, -2.0, 0.0, 3.0], None, 4.73818),
    (None, [1.0, 0.0, 1.0, 3.0], 4.64203),
    ([0.0, -2.0, 0.0, 3.0], [1.0, 0.0, 1.0, 3.0], 4.60829),
], ids=["active_primal_subspace", "active_dual_subspace", "both"])
def test_sigma_max_on_active_subspaces(x, y, want):
    v, iters = ob.Problem(fixtures.make_test_lp(), 2, 2).sigma_max_at(x, y, 0.01, 0.001, 1)
    assert abs(v - want) < 0.01 and iters < 300
    # and against the singular values of the projected matrix itself
    K = fixtures.make_test_lp().constraint_matrix.toarray()
    if x is not None:
        K = K[:, [0, 2, 3]]
    if y is not None:
        K = K[[0, 2, 3], :]
    assert abs(v - np.linalg.svd(K, compute_uv=False)[0]) < 0.01 * want


# ---- sharded_quadratic_program_test.cc: the data transformations of ShardedQuadraticProgram ----

def test_rescale_problem_basic():  # RescaleProblem.BasicTest, :91-123
    """TestDiagonalQp1 rescaled with rows [0.5] and columns [1, 0.5]: min 4 x1^2 + 0.25 x2^2 - x1 - 0.5 x2 + 5,
    -inf <= 0.5 x1 + 0.25 x2 <= 0.5, 1 <= x1 <= 2, -4 <= x2 <= 8; both copies of K carry the scaling."""
    p = ob.Problem(fixtures.diagonal_qp1(), num_threads=2, num_shards=10)
    p.rescale_with([1, 0.5], [0.5])
    got = p.get()
    assert got["clb"].tolist() == [-np.inf] and got["cub"].tolist() == [0.5]
    assert got["vlb"].tolist() == [1, -4] and got["vub"].tolist() == [2, 8]
    assert got["c"].tolist() == [-1, -0.5]
    assert got["val"].tolist() == [0.5, 0.25]            # K = [[0.5, 0.25]] by columns
    assert p.transposed_values().tolist() == [0.5, 0.25]  # K^T = [[0.5], [0.25]]
    assert got["q"].tolist() == [4, 0.25]


def test_replace_large_constraint_bounds_with_infinity():  # :125-135
    p = ob.Problem(fixtures.make_test_lp(), num_threads=2, num_shards=2)
    p.replace_large_constraint_bounds_with_infinity(3.0)
    got = p.get()
    assert got["clb"].tolist() == [np.inf, -np.inf, -np.inf, -1]
    assert got["cub"].tolist() == [np.inf, np.inf, np.inf, 1]


def test_sharded_quadratic_program_basic():  # ShardedQuadraticProgramTest.BasicTest, :32-49
    qp = fixtures.diagonal_qp1()
    p = ob.Problem(qp, num_threads=2, num_shards=10)
    assert (p.n, p.m) == (2, 1)
    assert p.transposed_values().tolist() == [1, 1]
    # K^T v and K v through the two copies.
    assert p.matvec(0, [3.0]).tolist() == [3, 3] and p.matvec(1, [2.0, 5.0]).tolist() == [7]


# ---- sharded_optimization_utils_test.cc:41-131: ShardedWeightedAverage (the update every accepted step applies) ----

def test_weighted_average_simple():  # :41-59
    avg, has_weight, terms = ob.weighted_average([[4, 1], [1, 7]], [1.0, 2.0], num_shards=2)
    assert has_weight and terms == 2 and avg.tolist() == [2.0, 5.0]
    avg, has_weight, terms = ob.weighted_average([[4, 1], [1, 7]], [1.0, 2.0], num_shards=2, clear_after_add=2)
    assert not has_weight and terms == 0


def test_weighted_average_of_nothing_is_zero():  # ZeroAverage, :91-99
    avg, has_weight, terms = ob.weighted_average(np.zeros((1, 1)), [])
    assert not has_weight and avg.tolist() == [0.0]


def test_weighted_average_of_one_vector_has_no_roundoff():  # AveragesEqualWithoutRoundoff, :103-115
    data = [1.0, 1.0 / 3, 3.0 / 7, 3.14159]
    for count in (1, 2, 3):
        avg, _, _ = ob.weighted_average([data] * count, [341.45, 1.4134, 7.23][:count])
        assert avg.tolist() == data  # exactly


def test_weighted_average_counts_zero_weight_terms():  # AddsZeroWeight, :117-127
    avg, has_weight, terms = ob.weighted_average([[1.0]], [0.0])
    assert not has_weight and avg.tolist() == [0.0] and terms == 1


# ---- sharded_optimization_utils_test.cc:596-612: projections onto the variable bounds / dual sign constraints ----

def test_projections_onto_bounds_of_test_lp():
    p = ob.Problem(fixtures.make_test_lp(), num_threads=2, num_shards=2)
    x, _ = p.project(x=[-3, -3, 5, 5])
    assert x.tolist() == [-3, -2, 5, 3.5]
    _, y = p.project(y=[1, 1, -1, -1])
    assert y.tolist() == [1, 0, 0, -1]


# ---- sharder_test.cc:260-405: the sharded BLAS-1 every vector pass of the reference goes through ----

@pytest.mark.parametrize("op, a, b, scalar, want", [
    ("SetZero", [1, 7, 0], None, 0.0, [0.0, 0.0, 0.0]),                         # :260 (the reference resizes a 2-vector to 3)
    ("ZeroVector", None, None, 0.0, [0.0, 0.0, 0.0]),                           # :267
    ("OnesVector", None, None, 0.0, [1.0, 1.0, 1.0]),                           # :272
    ("AddScaledVector", [4, 5, 20], [1, 7, 3], 2.0, [6, 19, 26]),              # :277
    ("AssignVector", [1, 7, 3], None, 0.0, [1, 7, 3]),                          # :285
    ("CloneVector", [1, 7, 3], None, 0.0, [1, 7, 3]),                           # :293
    ("CoefficientWiseProductInPlace", [4, 5, 20], [1, 2, 3], 0.0, [4, 10, 60]),  # :299
    ("CoefficientWiseQuotientInPlace", [4, 6, 20], [1, 2, 5], 0.0, [4, 3, 4]),  # :308
])
def test_sharded_vector_operations(op, a, b, scalar, want):
    _, got = ob.blas1(op, 3, a, b, scalar)
    assert got.tolist() == want


@pytest.mark.parametrize("op, a, b, want, tol", [
    ("Dot", [1, 2, 3], [4, 5, 6], 32.0, 1e-13),                     # :317
    ("LInfNorm", [-1, 2, -3], None, 3.0, 0.0),                      # :325
    ("L1Norm", [-1, 2, -3], None, 6.0, 0.0),                        # :339
    ("SquaredNorm", [1, 2, 3], None, 14.0, 1e-13),                  # :353
    ("Norm", [1, 2, 3], None, 14.0 ** 0.5, 1e-13),                  # :360
    ("SquaredDistance", [1, 1, 1], [1, 2, 3], 5.0, 1e-13),          # :367
    ("Distance", [1, 1, 1], [1, 2, 3], 5.0 ** 0.5, 1e-13),          # :375
    ("ScaledLInfNorm", [-1, 2, -3], [4, 6, 1], 12.0, 0.0),          # :383
    ("ScaledSquaredNorm", [-1, 2, -3], [4, 6, 1], 169.0, 0.0),      # :391
    ("ScaledNorm", [-1, 2, -3], [4, 6, 1], 13.0, 0.0),              # :399
])
def test_sharded_reductions(op, a, b, want, tol):
    got, _ = ob.blas1(op, 3, a, b)
    assert abs(got - want) <= tol


def test_norms_of_empty_vectors():  # LInfNormTest.EmptyExample :332, L1NormTest.EmptyExample :346
    assert ob.blas1("LInfNorm", 0, np.zeros(0))[0] == 0 and ob.blas1("L1Norm", 0, np.zeros(0))[0] == 0


@pytest.mark.parametrize("size", [10, 1000, 100 * 1000])
def test_various_sizes_threaded(size):  # VariousSizesTest.LargeMatVec / LargeVectors, sharder_test.cc:431-460: 5 threads, 15 shards
    import scipy.sparse as sp
    rng = np.random.default_rng(size)
    nnz = 8 * size
    K = sp.coo_matrix((rng.standard_normal(nnz), (rng.integers(0, size, nnz), rng.integers(0, size, nnz))), shape=(size, size)).tocsc()
    K.sum_duplicates()
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(size, size)
    qp.constraint_matrix = K
    p = ob.Problem(qp, num_threads=5, num_shards=15)
    v = rng.uniform(-1, 1, size)
    assert np.linalg.norm(K.T @ v - p.matvec(0, v)) <= 1.0e-8
    assert np.linalg.norm(K @ v - p.matvec(1, v)) <= 1.0e-8
    got, _ = ob.blas1("SquaredNorm", size, v, num_shards=5)
    assert abs(got - float(v @ v)) <= size * 1.0e-14


# ---- iteration_stats_test.cc:132-161: random projections (values unpinned: absl::Gaussian; the properties hold) ----

def test_random_projection_of_the_zero_vector_is_zero():
    p = ob.Problem(fixtures.make_test_lp(), num_threads=2, num_shards=10)
    primal, dual = p.random_projections(np.zeros(4), np.zeros(4), [1])
    assert primal.tolist() == [0.0] and dual.tolist() == [0.0]


def test_two_random_projections_of_a_vector():
    p = ob.Problem(fixtures.make_test_lp(), num_threads=2, num_shards=10)
    primal, dual = p.random_projections(np.ones(4), np.zeros(4), [1, 2])
    assert len(primal) == 2 and len(dual) == 2
    # The primal point has norm 2; a projection onto a unit direction can only reduce it, and is 0 with probability 0.
    assert all(-2.0 <= v <= 2.0 and v != 0.0 for v in primal)
    assert all(v == 0.0 for v in dual)


# ---- trust_region_test.cc:758-792: ComputeLocalizedLagrangianBounds with the caller's K x and K^T y ----

@pytest.mark.parametrize("max_norm", [False, True], ids=["euclidean", "max"])
@pytest.mark.parametrize("diag", [False, True], ids=["linear_time", "diagonal_qp"])
def test_localized_bounds_accept_cached_products(max_norm, diag):
    p = ob.Problem(fixtures.make_test_lp(), num_threads=2, num_shards=2)
    x, y = [0.0, 0.0, 0.0, 3.0], np.zeros(4)  # x_3 has a lower bound of 2.5
    primal2 = 0.5 * (1.0 + 64.0 + 1.0 + 0.25)
    dual2 = 0.5 * (4.0 + 2.375 * 2.375 + 4.0 / 9.0)
    radius = math.sqrt(max(primal2, dual2)) if max_norm else math.sqrt(primal2 + dual2)
    b = p.localized_bounds_cached(x, y, [6.0, 0.0, 0.0, -3.0], np.zeros(4), 1.0, radius, use_diagonal=diag, tol=1e-6, max_norm=max_norm)
    assert b["lagrangian_value"] == 3.0
    assert b["lower_bound"] <= -20.0 <= b["upper_bound"]
    # and the cached products change nothing when they are the true ones
    same = p.localized_bounds(x, y, 0 if max_norm else 1, 1.0, radius, diag, 1e-6)
    assert (b["lower_bound"], b["upper_bound"]) == (same["lower_bound"], same["upper_bound"])
