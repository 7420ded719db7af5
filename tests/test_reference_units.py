"""Known-answer tests re-typed from the reference's unit tests, run against BOTH the CPU
oracle (pins the oracle, `-m "not gpu"`) and the CUDA kernels through the C ABI (`-m gpu`).

Sources: ortools/pdlp/sharder_test.cc:251-428, sharded_optimization_utils_test.cc:130-400,
iteration_stats_test.cc:44-254, trust_region_test.cc:615-960,1071-1090.
"""
import math

import numpy as np
import pytest
import scipy.sparse as sp

import fixtures
import pdlp_b200

P = pdlp_b200.PrimalDualHybridGradientParams
INF = np.inf


def small_matrix_qp():
    # sharder_test.cc:41-50: [[7, -0.5, 0, 0], [1, 0, 3, 2], [-1, 0, 0, 5]]
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(4, 3)
    qp.constraint_matrix = sp.csc_matrix(np.array([[7, -0.5, 0, 0], [1, 0, 3, 2], [-1, 0, 0, 5.0]]))
    return qp


def eq(a, b):
    """EXPECT_DOUBLE_EQ: within 4 ulp."""
    a, b = float(a), float(b)
    if a == b:
        return True
    return abs(a - b) <= 4 * np.spacing(max(abs(a), abs(b)))


def assert_all_eq(got, want):
    got, want = np.asarray(got, dtype=float), np.asarray(want, dtype=float)
    assert got.shape == want.shape, (got, want)
    for g, w in zip(got, want):
        assert eq(g, w), (got, want)


# ---- sharder_test.cc ----
def test_transposed_matrix_vector_product(backend):
    p = backend.problem(small_matrix_qp())
    assert_all_eq(p.matvec(0, [1, 2, 3.0]), [6, -0.5, 6, 19])  # :251-258
    assert_all_eq(p.matvec(1, [1, 2, 3, 4.0]), [6, 18, 19])


def test_scaled_col_norms(backend):
    p = backend.problem(small_matrix_qp())
    rs, cs = [1, -2, 1.0], [1, 2, -1, -1.0]
    assert_all_eq(p.col_norms(False, rs, cs), [7, 1, 6, 5])  # :407-416
    assert_all_eq(p.col_norms(True, rs, cs), [math.sqrt(54), 1, 6, math.sqrt(41)])  # :418-428


# ---- sharded_optimization_utils_test.cc: ComputeStats ----
STATS = {
    "test_lp": (fixtures.make_test_lp, dict(
        num_variables=4, num_constraints=4, constraint_matrix_col_min_l_inf_norm=1.0, constraint_matrix_row_min_l_inf_norm=1.0,
        constraint_matrix_num_nonzeros=9, constraint_matrix_abs_max=4.0, constraint_matrix_abs_min=1.0, constraint_matrix_abs_avg=14.5 / 9.0,
        constraint_matrix_l2_norm=math.sqrt(31.25), objective_vector_abs_max=5.5, objective_vector_abs_min=1.0, objective_vector_abs_avg=2.375,
        objective_vector_l2_norm=math.sqrt(36.25), objective_matrix_num_nonzeros=0, objective_matrix_abs_max=0.0, objective_matrix_abs_min=0.0,
        objective_matrix_abs_avg=math.nan, objective_matrix_l2_norm=0.0, variable_bound_gaps_num_finite=1, variable_bound_gaps_max=1.0,
        variable_bound_gaps_min=1.0, variable_bound_gaps_avg=1.0, variable_bound_gaps_l2_norm=1.0, combined_bounds_max=12.0, combined_bounds_min=1.0,
        combined_bounds_avg=6.0, combined_bounds_l2_norm=math.sqrt(210.0))),
    "tiny_lp": (fixtures.tiny_lp, dict(
        num_variables=4, num_constraints=3, constraint_matrix_col_min_l_inf_norm=1.0, constraint_matrix_row_min_l_inf_norm=1.0,
        constraint_matrix_num_nonzeros=8, constraint_matrix_abs_max=2.0, constraint_matrix_abs_min=1.0, constraint_matrix_abs_avg=1.25,
        constraint_matrix_l2_norm=math.sqrt(14.0), objective_vector_abs_max=5.0, objective_vector_abs_min=1.0, objective_vector_abs_avg=2.25,
        objective_vector_l2_norm=math.sqrt(31.0), objective_matrix_num_nonzeros=0, objective_matrix_abs_max=0.0, objective_matrix_abs_min=0.0,
        objective_matrix_abs_avg=math.nan, objective_matrix_l2_norm=0.0, variable_bound_gaps_num_finite=4, variable_bound_gaps_max=6.0,
        variable_bound_gaps_min=2.0, variable_bound_gaps_avg=3.75, variable_bound_gaps_l2_norm=math.sqrt(65.0), combined_bounds_max=12.0,
        combined_bounds_min=1.0, combined_bounds_avg=20.0 / 3.0, combined_bounds_l2_norm=math.sqrt(194.0))),
    "diagonal_qp1": (fixtures.diagonal_qp1, dict(
        num_variables=2, num_constraints=1, constraint_matrix_col_min_l_inf_norm=1.0, constraint_matrix_row_min_l_inf_norm=1.0,
        constraint_matrix_num_nonzeros=2, constraint_matrix_abs_max=1.0, constraint_matrix_abs_min=1.0, constraint_matrix_abs_avg=1.0,
        constraint_matrix_l2_norm=math.sqrt(2.0), objective_vector_abs_max=1.0, objective_vector_abs_min=1.0, objective_vector_abs_avg=1.0,
        objective_vector_l2_norm=math.sqrt(2.0), objective_matrix_num_nonzeros=2, objective_matrix_abs_max=4.0, objective_matrix_abs_min=1.0,
        objective_matrix_abs_avg=2.5, objective_matrix_l2_norm=math.sqrt(17.0), variable_bound_gaps_num_finite=2, variable_bound_gaps_max=6.0,
        variable_bound_gaps_min=1.0, variable_bound_gaps_avg=3.5, variable_bound_gaps_l2_norm=math.sqrt(37.0), combined_bounds_max=1.0,
        combined_bounds_min=1.0, combined_bounds_avg=1.0, combined_bounds_l2_norm=1.0)),
    "no_finite_gaps": (fixtures.small_invalid_problem_lp, dict(
        variable_bound_gaps_num_finite=0, variable_bound_gaps_max=0.0, variable_bound_gaps_min=0.0, variable_bound_gaps_avg=math.nan,
        variable_bound_gaps_l2_norm=0.0)),
    "lp_without_constraints": (fixtures.lp_without_constraints, dict(
        constraint_matrix_num_nonzeros=0, constraint_matrix_abs_max=0.0, constraint_matrix_abs_min=0.0, constraint_matrix_abs_avg=math.nan,
        constraint_matrix_l2_norm=0.0, constraint_matrix_col_min_l_inf_norm=0.0, constraint_matrix_row_min_l_inf_norm=0.0, combined_bounds_max=0.0,
        combined_bounds_min=0.0, combined_bounds_avg=math.nan, combined_bounds_l2_norm=0.0)),
}


def modified_qp1():
    qp = fixtures.diagonal_qp1()
    qp.set_objective_matrix_diagonal([2.0, 0.0])
    return qp


def empty_qp():
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(0, 0)
    return qp


STATS["modified_qp1"] = (modified_qp1, dict(objective_matrix_num_nonzeros=1, objective_matrix_abs_max=2.0, objective_matrix_abs_min=2.0,
                                            objective_matrix_abs_avg=1.0, objective_matrix_l2_norm=2.0))
STATS["empty"] = (empty_qp, dict(
    num_variables=0, num_constraints=0, constraint_matrix_col_min_l_inf_norm=0.0, constraint_matrix_row_min_l_inf_norm=0.0, constraint_matrix_num_nonzeros=0,
    constraint_matrix_abs_max=0.0, constraint_matrix_abs_min=0.0, constraint_matrix_abs_avg=math.nan, constraint_matrix_l2_norm=0.0,
    objective_vector_abs_max=0.0, objective_vector_abs_min=0.0, objective_vector_abs_avg=math.nan, objective_vector_l2_norm=0.0,
    objective_matrix_num_nonzeros=0, objective_matrix_abs_max=0.0, objective_matrix_abs_min=0.0, objective_matrix_abs_avg=math.nan,
    objective_matrix_l2_norm=0.0, variable_bound_gaps_num_finite=0, variable_bound_gaps_max=0.0, variable_bound_gaps_min=0.0,
    variable_bound_gaps_avg=math.nan, variable_bound_gaps_l2_norm=0.0, combined_bounds_max=0.0, combined_bounds_min=0.0, combined_bounds_avg=math.nan,
    combined_bounds_l2_norm=0.0))


@pytest.mark.parametrize("case", sorted(STATS))
def test_problem_stats(backend, case):
    make, want = STATS[case]
    stats = backend.problem(make()).stats()
    for field, value in want.items():
        got = getattr(stats, field)
        if isinstance(value, float) and math.isnan(value):
            assert math.isnan(got), field
        else:
            assert eq(got, value), (field, got, value)


# ---- rescaling (sharded_optimization_utils_test.cc:373-400) ----
def test_apply_rescaling_test_lp(backend):
    rs, cs, vals = backend.problem(fixtures.make_test_lp()).rescale(1, True)
    np.testing.assert_allclose(rs, [1 / math.sqrt(2.0 * 1.5275), 1 / math.sqrt(1.0 * 0.9574), 1 / math.sqrt(4.0 * 1.0), 1 / math.sqrt(1.5 * 1.1547)], atol=1e-4)
    np.testing.assert_allclose(cs, [1 / math.sqrt(4.0 * 1.3229), 1 / math.sqrt(1.0 * 0.7071), 1 / math.sqrt(1.5 * 1.4142), 1 / math.sqrt(2.0 * 1.1547)], atol=1e-4)
    # the scaled matrix is diag(rs) K diag(cs)
    K = fixtures.make_test_lp().constraint_matrix.tocsc()
    want = sp.diags(rs) @ K @ sp.diags(cs)
    np.testing.assert_allclose(vals, want.tocsc().data, rtol=1e-14)


def test_ruiz_converges_to_unit_norms(backend):
    # LInfRuizRescaling.Convergence (:359-371)
    qp = fixtures.make_test_lp()
    rs, cs, vals = backend.problem(qp).rescale(20, False)
    K = (sp.diags(rs) @ qp.constraint_matrix @ sp.diags(cs)).toarray()
    np.testing.assert_allclose(np.abs(K).max(axis=0), np.ones(4), atol=1e-4)
    np.testing.assert_allclose(np.abs(K).max(axis=1), np.ones(4), atol=1e-4)


# ---- iteration_stats_test.cc ----
def conv(backend, qp, x, y, flag=None):
    p = P()
    if flag is not None:
        p.handle_some_primal_gradients_on_finite_bounds_as_residuals = flag
    return backend.problem(qp).kkt_stats(p, np.array(x, dtype=float), np.array(y, dtype=float)).convergence_information[0]


def test_corrected_dual_simple_lp_with_suboptimal_dual(backend):
    c = conv(backend, fixtures.make_test_lp(), [0, 0, 6, 2.5], [-2, 0, 2.375, 1])
    assert eq(c.dual_objective, -36.5) and eq(c.corrected_dual_objective, -36.5)


def test_corrected_dual_far_from_bound_as_residuals(backend):
    c = conv(backend, fixtures.make_test_lp(), [0, 0, 2, 2.5], [-2, 0, 2.375, 1], flag=True)
    assert eq(c.dual_objective, -33.5) and eq(c.corrected_dual_objective, -36.5)
    assert eq(c.l_inf_dual_residual, 0.5) and eq(c.l2_dual_residual, 0.5) and eq(c.l_inf_componentwise_dual_residual, 0.25)


def test_corrected_dual_far_from_bound_as_reduced_costs(backend):
    c = conv(backend, fixtures.make_test_lp(), [0, 0, 2, 2.5], [-2, 0, 2.375, 1], flag=False)
    assert eq(c.dual_objective, -36.5) and eq(c.corrected_dual_objective, -36.5)
    assert c.l_inf_dual_residual == 0.0 and c.l2_dual_residual == 0.0 and c.l_inf_componentwise_dual_residual == 0.0


def test_corrected_dual_qp_suboptimal(backend):
    c = conv(backend, fixtures.diagonal_qp1(), [-2.0, 2.0], [-3])
    assert eq(c.corrected_dual_objective, -28.0)


def test_convergence_information_at_test_lp_optimum(backend):
    # the optimum of TestLp (test_util.h:45-48): zero residuals, both objectives -34
    c = conv(backend, fixtures.make_test_lp(), [-1, 8, 1, 2.5], [-2, 0, 2.375, 2.0 / 3])
    assert abs(c.primal_objective + 34) < 1e-12 and abs(c.dual_objective + 34) < 1e-12
    assert c.l_inf_primal_residual < 1e-12 and c.l_inf_dual_residual < 1e-12
    assert eq(c.l_inf_primal_variable, 8.0) and eq(c.l_inf_dual_variable, 2.375)


def test_reduced_costs_simple_lp(backend):
    p = backend.problem(fixtures.make_test_lp())
    x, y = [0.0, -2.0, 6.0, 3.5], [1.0, 0.0, 0.0, -2.0]
    assert_all_eq(p.reduced_costs(P(), x, y), [0, 0, 0, -3.0])
    assert_all_eq(p.reduced_costs(P(), x, y, True), [0, 0, 0, -4.0])


def test_reduced_costs_with_gap_residuals(backend):
    p = backend.problem(fixtures.make_test_lp())
    t, f = P(), P()
    t.handle_some_primal_gradients_on_finite_bounds_as_residuals = True
    f.handle_some_primal_gradients_on_finite_bounds_as_residuals = False
    y = [1.0, 0.0, 0.0, -1.0]
    assert_all_eq(p.reduced_costs(t, np.zeros(4), y), [0, 0, 0, 0])
    assert_all_eq(p.reduced_costs(f, np.zeros(4), y), [0, 0, -0.5, -2.0])
    x2 = [0.0, 0.0, 4.0, 3.0]
    assert_all_eq(p.reduced_costs(t, x2, y), [0, 0, -0.5, -2.0])
    assert_all_eq(p.reduced_costs(f, x2, y), [0, 0, -0.5, -2.0])


def test_reduced_costs_simple_qp(backend):
    p = backend.problem(fixtures.diagonal_qp1())
    t, f = P(), P()
    t.handle_some_primal_gradients_on_finite_bounds_as_residuals = True
    f.handle_some_primal_gradients_on_finite_bounds_as_residuals = False
    x, y = [1.0, 2.0], [0.0]
    assert_all_eq(p.reduced_costs(t, x, y), [3.0, 0.0])
    assert_all_eq(p.reduced_costs(f, x, y), [3.0, 1.0])
    assert_all_eq(p.reduced_costs(t, x, y, True), [0.0, 0.0])
    assert_all_eq(p.reduced_costs(f, x, y, True), [0.0, 0.0])


# ---- trust_region_test.cc: ComputeLocalizedLagrangianBounds (Euclidean norm, the one PDHG uses) ----
def one_dim(q=None):
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(1, 1)
    qp.constraint_matrix = sp.csc_matrix(np.array([[1.0]]))
    qp.constraint_lower_bounds = np.array([0.0])
    qp.constraint_upper_bounds = np.array([1.0])
    qp.objective_vector = np.array([1.0])
    if q is not None:
        qp.set_objective_matrix_diagonal([q])
    return qp


@pytest.mark.parametrize("diag", [False, True])
def test_localized_bounds_zero_gap_at_optimal(backend, diag):
    b = backend.problem(fixtures.make_test_lp()).localized_bounds([-1.0, 8.0, 1.0, 2.5], [-2.0, 0.0, 2.375, 2.0 / 3.0], 1.0, 1.0, diag, 1e-2)
    assert eq(b["radius"], 1.0) and eq(b["lagrangian_value"], -20.0)
    tol = 1e-9 if diag else 0.0
    assert abs(b["lower_bound"] + 20.0) <= tol + 1e-13 and abs(b["upper_bound"] + 20.0) <= tol + 1e-13


@pytest.mark.parametrize("diag", [False, True])
def test_localized_bounds_optimal_in_range(backend, diag):
    r = math.sqrt(0.5 * (1.0 + 64.0 + 1.0 + 0.25) + 0.5 * (4.0 + 2.375 * 2.375 + 4.0 / 9.0))
    b = backend.problem(fixtures.make_test_lp()).localized_bounds([0, 0, 0, 3.0], np.zeros(4), 1.0, r, diag, 1e-6)
    assert eq(b["lagrangian_value"], 3.0)
    assert b["lower_bound"] <= -20.0 <= b["upper_bound"]


@pytest.mark.parametrize("diag", [False, True])
def test_localized_bounds_optimal_not_in_range(backend, diag):
    b = backend.problem(fixtures.make_test_lp()).localized_bounds([0, 0, 0, 3.0], np.zeros(4), 1.0, 0.1, diag, 1e-6)
    assert eq(b["lagrangian_value"], 3.0)
    assert abs(b["lower_bound"] - (3.0 - 0.1 * math.sqrt(2) * 36.25 / math.sqrt(76.25))) < 1e-6
    assert abs(b["upper_bound"] - (3.0 + 0.1 * math.sqrt(2) * 40 / math.sqrt(76.25))) < 1e-6


@pytest.mark.parametrize("diag", [False, True])
def test_localized_bounds_one_dim_lp(backend, diag):
    b = backend.problem(one_dim()).localized_bounds([0.0], [-1.0], 1.0, 1.0 / math.sqrt(2.0), diag, 1e-6)
    assert eq(b["lagrangian_value"], -1.0)
    tol = 1e-6 if diag else 1e-15
    assert abs(b["lower_bound"] - (-1.0 - 4.0 / math.sqrt(5))) <= tol
    assert abs(b["upper_bound"] - (-1.0 + 1.0 / math.sqrt(5))) <= tol


@pytest.mark.parametrize("diag", [False, True])
def test_localized_bounds_one_dim_lp_large_primal_weight(backend, diag):
    b = backend.problem(one_dim()).localized_bounds([0.0], [-1.0], 100.0, 1.0 / math.sqrt(2.0), diag, 1e-8)
    assert eq(b["lagrangian_value"], -1.0)
    tol = 10.002e-8 if diag else 1e-14
    assert abs(b["upper_bound"] - b["lower_bound"] - 10.00199980003999) <= tol


def test_localized_bounds_one_dim_qp_unit_weight(backend):
    # SolvesForTestQpUnitWeight (:1071-1090)
    b = backend.problem(one_dim(q=2.0)).localized_bounds([0.0], [-1.0], 1.0, 0.5, True, 1e-6)
    assert eq(b["lagrangian_value"], -1.0)
    assert abs(b["upper_bound"] - (-1.0 + 0.5)) < 1e-5
    assert abs(b["lower_bound"] - (-1.0 - 0.75)) < 1e-5


# kMaxNorm variants (NormsBehaveDifferently, :756-805; on the device through pdlp_b200_localized_bounds_max_norm)
def test_localized_bounds_max_norm(backend):
    p = backend.problem(fixtures.make_test_lp())
    b = p.localized_bounds([0, 0, 0, 3.0], np.zeros(4), 1.0, 0.1, max_norm=True)
    assert abs(b["lower_bound"] - (3.0 - 0.1 * math.sqrt(2) * math.sqrt(36.25))) < 1e-6
    assert abs(b["upper_bound"] - (3.0 + 0.1 * math.sqrt(2) * math.sqrt(40.0))) < 1e-6
    # ProcessesPrimalWeight (:730-754)
    b = p.localized_bounds([0, 0, 0, 3.0], np.zeros(4), 100.0, 0.1, max_norm=True)
    assert 3.0 - 0.28 <= b["lower_bound"] <= 3.0 - 0.028
    assert 3.0 + 2.8 <= b["upper_bound"] <= 3.0 + 28
    q = backend.problem(one_dim())
    b = q.localized_bounds([0.0], [-1.0], 1.0, 1.0 / math.sqrt(2.0), max_norm=True)
    assert eq(b["lower_bound"], -3.0) and eq(b["upper_bound"], 0.0)
    b = q.localized_bounds([0.0], [-1.0], 100.0, 1.0 / math.sqrt(2.0), max_norm=True)
    assert eq(b["lower_bound"], -1.2) and eq(b["upper_bound"], 9.0)
