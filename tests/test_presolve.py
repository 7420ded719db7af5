"""The presolve hand-off (include/pdlp_b200.h: pdlp_b200_solve_presolved), the device path's form of the
reference's Glop branch (primal_dual_hybrid_gradient.cc:1145-1202 ApplyPresolveIfEnabled, :986-1034 the
"presolver solved it" result, :1288-1341 RecoverOriginalSolution, :1530-1548 statistics on the original
problem).  Glop is not available here, so the tests bring a small presolver of their own - fixed variables
substituted out, constraints that become empty dropped - which is enough to exercise every rule of the
hand-off: sizes of the two statistics blocks, statistics and termination on the ORIGINAL problem, the
recovered solution and its reduced costs, the ignored warm start, kPresolveTermination."""
import numpy as np
import pytest
import scipy.sparse as sp

import fixtures
import oracle_binding as ob
import pdlp_b200

TR = pdlp_b200.TerminationReason
PT = pdlp_b200.PointType


def toy_presolve(qp):
    """Removes fixed variables (lb == ub) and the constraints that are empty afterwards.
    Returns (presolved qp or None, postsolve, verdict): verdict is None when something is left to solve,
    else the TerminationReason the presolver claims."""
    K = sp.csc_matrix(qp.constraint_matrix)
    lb, ub = qp.variable_lower_bounds, qp.variable_upper_bounds
    fixed = lb == ub
    keep = ~fixed
    x_fixed = np.where(fixed, lb, 0.0)
    shift = K @ x_fixed
    Kk = K[:, keep]
    row_keep = np.diff(sp.csr_matrix(Kk).indptr) > 0
    clb, cub = qp.constraint_lower_bounds - shift, qp.constraint_upper_bounds - shift
    n, m = len(lb), K.shape[0]

    def postsolve(x, y):
        xo = x_fixed.copy()
        xo[keep] = x
        yo = np.zeros(m)
        yo[row_keep] = y
        return xo, yo

    if not keep.any():
        feasible = np.all(clb <= 1e-12) and np.all(cub >= -1e-12)
        return None, postsolve, (TR.TERMINATION_REASON_OPTIMAL if feasible else TR.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE)
    pre = pdlp_b200.QuadraticProgram()
    pre.constraint_matrix = sp.csc_matrix(Kk[row_keep, :])
    pre.objective_vector = qp.objective_vector[keep]
    pre.variable_lower_bounds, pre.variable_upper_bounds = lb[keep], ub[keep]
    pre.constraint_lower_bounds, pre.constraint_upper_bounds = clb[row_keep], cub[row_keep]
    pre.objective_offset = qp.objective_offset + float(qp.objective_vector[fixed] @ lb[fixed])
    pre.objective_scaling_factor = qp.objective_scaling_factor
    return pre, postsolve, None


def lp_with_fixed_variables(seed=3):
    qp = fixtures.random_lp(120, 260, 4, seed)
    rng = np.random.default_rng(seed)
    fix = rng.random(260) < 0.3
    value = np.clip(getattr(qp, "planted_primal"), qp.variable_lower_bounds, qp.variable_upper_bounds)
    qp.variable_lower_bounds = np.where(fix, value, qp.variable_lower_bounds)
    qp.variable_upper_bounds = np.where(fix, value, qp.variable_upper_bounds)
    return qp


def params(eps=1e-7):
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps
    return p


def test_the_toy_presolver_keeps_the_optimum():  # the test's own tool, checked against the oracle on CPU
    qp = lp_with_fixed_variables()
    pre, postsolve, verdict = toy_presolve(qp)
    assert verdict is None and pre.constraint_matrix.shape[1] < qp.constraint_matrix.shape[1]
    a, b = ob.solve(qp, params()), ob.solve(pre, params())
    obj = lambda r: [c for c in r.solve_log_dict["solution_stats"]["convergence_information"] if c["candidate_type"] == r.solve_log_dict["solution_type"]][0]["primal_objective"]
    assert abs(obj(a) - obj(b)) <= 1e-5 * max(1.0, abs(obj(a)))
    x, y = postsolve(b.primal_solution, b.dual_solution)
    assert abs(float(qp.objective_vector @ x) + qp.objective_offset - obj(a)) <= 1e-5 * max(1.0, abs(obj(a)))


def test_bad_arguments_are_rejected_before_any_device_work():
    qp = lp_with_fixed_variables()
    pre, postsolve, _ = toy_presolve(qp)
    pre.constraint_lower_bounds = pre.constraint_lower_bounds[:-1]
    with pytest.raises(ValueError, match="presolved problem"):
        pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, pre, postsolve, params())


@pytest.mark.gpu
def test_presolved_solve_reports_on_the_original_problem():
    qp = lp_with_fixed_variables()
    pre, postsolve, _ = toy_presolve(qp)
    n, m = qp.constraint_matrix.shape[1], qp.constraint_matrix.shape[0]
    types = []
    warm = pdlp_b200.PrimalAndDualSolution(np.ones(n), np.zeros(m))  # ignored under presolve (:1164-1168)
    got = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, pre, postsolve, params(), initial_solution=warm,
                                                          iteration_stats_callback=lambda info: types.append(info.iteration_type))
    direct = pdlp_b200.primal_dual_hybrid_gradient(qp, params())
    log = got.solve_log
    assert log.termination_reason == direct.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    # sizes: the original in original_problem_stats, the presolved problem in preprocessed_problem_stats (pdhg_test.cc:311-314 style)
    assert log.original_problem_stats.num_variables == n and log.original_problem_stats.num_constraints == m
    assert log.preprocessed_problem_stats.num_variables == pre.constraint_matrix.shape[1] < n
    assert log.preprocessed_problem_stats.num_constraints == pre.constraint_matrix.shape[0] <= m
    assert got.primal_solution.size == n and got.dual_solution.size == m and got.reduced_costs.size == n
    info = lambda r: [c for c in r.solve_log.solution_stats.convergence_information if c.candidate_type == r.solve_log.solution_type][0]
    assert abs(info(got).primal_objective - info(direct).primal_objective) <= 1e-5 * max(1.0, abs(info(direct).primal_objective))
    fixed = qp.variable_lower_bounds == qp.variable_upper_bounds
    np.testing.assert_array_equal(got.primal_solution[fixed], qp.variable_lower_bounds[fixed])
    # the statistics in the log are those of the ORIGINAL problem at the recovered point: the oracle agrees
    stats = ob.Problem(qp).convergence_info(params(), np.ones(n), np.ones(m), got.primal_solution, got.dual_solution)
    for key in ("primal_objective", "dual_objective", "l2_primal_residual", "l2_dual_residual"):
        assert abs(stats[key] - getattr(info(got), key)) <= 1e-9 * max(1.0, abs(stats[key])), key
    st = ob.Problem(qp).stats()
    assert stats["l2_primal_residual"] <= 1.0001e-7 * (1 + st["combined_bounds_l2_norm"])
    assert stats["l2_dual_residual"] <= 1.0001e-7 * (1 + st["objective_vector_l2_norm"])
    # reduced costs with respect to the original problem (:1578-1581)
    rc = ob.Problem(qp).reduced_costs(params(), got.primal_solution, got.dual_solution)
    np.testing.assert_allclose(got.reduced_costs, rc, rtol=1e-9, atol=1e-9)
    assert types and types[-1] == 4 and all(t == 0 for t in types[:-1])  # kNormal ..., kNormalTermination


@pytest.mark.gpu
def test_presolver_that_solves_everything_gives_a_presolver_solution():
    qp = fixtures.make_test_lp()
    x_opt = np.array([-1.0, 8.0, 1.0, 2.5])
    qp.variable_lower_bounds = x_opt.copy()
    qp.variable_upper_bounds = x_opt.copy()
    pre, postsolve, verdict = toy_presolve(qp)
    assert pre is None and verdict == TR.TERMINATION_REASON_OPTIMAL
    types = []
    # The toy postsolve returns zero duals: the recovered point is primal feasible with the right objective, but not dual
    # optimal, so the claim "OPTIMAL" is not confirmed and the reference's rule downgrades it (:1016-1024).
    got = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, None, postsolve, params(), presolve_status=verdict,
                                                          iteration_stats_callback=lambda info: types.append(info.iteration_type))
    assert got.solve_log.solution_type == PT.POINT_TYPE_PRESOLVER_SOLUTION
    assert got.solve_log.termination_reason in (TR.TERMINATION_REASON_OPTIMAL, TR.TERMINATION_REASON_NUMERICAL_ERROR)
    assert got.solve_log.iteration_count == 0 and types == [3]  # kPresolveTermination
    np.testing.assert_array_equal(got.primal_solution, x_opt)
    c = got.solve_log.solution_stats.convergence_information[0]
    assert c.candidate_type == PT.POINT_TYPE_PRESOLVER_SOLUTION and abs(c.primal_objective + 34.0) < 1e-12
    # With the true duals the same claim is confirmed.
    y_opt = np.array([-2.0, 0.0, 2.375, 2.0 / 3.0])
    confirmed = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, None, lambda x, y: (x_opt, y_opt), params(1e-6), presolve_status=verdict)
    assert confirmed.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    # A presolver that reports infeasibility is passed through when the point earns nothing better.
    bad = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, None, postsolve, params(), presolve_status=TR.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE)
    assert bad.solve_log.termination_reason in (TR.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE, TR.TERMINATION_REASON_OPTIMAL)


@pytest.mark.gpu
def test_presolve_solves_to_optimality_goldens():  # PresolveTest.PresolveSolvesToOptimality, :1976-2003
    """Glop's preprocessor solves TestLp completely; here the caller's presolver does (it hands back the optimum).  The
    statistics of the result are the reference's goldens for that point."""
    x_opt, y_opt = np.array([-1.0, 8.0, 1.0, 2.5]), np.array([-2.0, 0.0, 2.375, 2.0 / 3.0])
    p = pdlp_b200.PrimalDualHybridGradientParams()
    out = pdlp_b200.primal_dual_hybrid_gradient_presolved(fixtures.make_test_lp(), None, lambda x, y: (x_opt, y_opt), p,
                                                          presolve_status=pdlp_b200.glop_status_to_termination_reason(pdlp_b200.GlopProblemStatus.OPTIMAL))
    log = out.solve_log
    assert log.termination_reason == TR.TERMINATION_REASON_OPTIMAL and log.iteration_count == 0
    assert log.solution_type == PT.POINT_TYPE_PRESOLVER_SOLUTION
    np.testing.assert_allclose(out.primal_solution, x_opt, atol=1e-10, rtol=0)
    np.testing.assert_allclose(out.dual_solution, y_opt, atol=1e-10, rtol=0)
    infos = [c for c in log.solution_stats.convergence_information if c.candidate_type == PT.POINT_TYPE_PRESOLVER_SOLUTION]
    assert len(infos) == 1
    c = infos[0]
    for name, want in (("primal_objective", -34.0), ("dual_objective", -34.0), ("corrected_dual_objective", -34.0), ("l_inf_primal_residual", 0.0),
                       ("l2_primal_residual", 0.0), ("l_inf_dual_residual", 0.0), ("l2_dual_residual", 0.0), ("l_inf_primal_variable", 8.0),
                       ("l2_primal_variable", 8.5), ("l_inf_dual_variable", 2.375)):
        assert abs(getattr(c, name) - want) <= 4e-15 * max(1.0, abs(want)), (name, getattr(c, name))  # EXPECT_DOUBLE_EQ: 4 ulps
    assert abs(c.l2_dual_variable - 3.17569983538) <= 1e-10


@pytest.mark.gpu
def test_presolve_infeasible():  # PresolveTest.PresolveInfeasible, :2027-2036
    qp = fixtures.small_primal_infeasible_lp()
    n, m = len(qp.objective_vector), len(qp.constraint_lower_bounds)
    verdict = pdlp_b200.glop_status_to_termination_reason(pdlp_b200.GlopProblemStatus.PRIMAL_INFEASIBLE)
    out = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, None, lambda x, y: (np.zeros(n), np.zeros(m)), pdlp_b200.PrimalDualHybridGradientParams(),
                                                          presolve_status=verdict)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE
    assert out.solve_log.solution_type == PT.POINT_TYPE_PRESOLVER_SOLUTION
    assert out.solve_log.iteration_count == 0


# ---- internal::ComputeStatuses (primal_dual_hybrid_gradient.h:164, .cc:2835-2883): host code, no GPU ----

CS, VS = pdlp_b200.ConstraintStatus, pdlp_b200.VariableStatus


def test_compute_statuses_at_optimum():  # primal_dual_hybrid_gradient_test.cc:2076-2092
    lp = fixtures.make_test_lp()
    constraint_statuses, variable_statuses = pdlp_b200.compute_statuses(lp, [-1, 8, 1, 2.5], [-2, 0, 2.375, 2.0 / 3])
    assert constraint_statuses == [CS.FIXED_VALUE, CS.BASIC, CS.AT_LOWER_BOUND, CS.AT_LOWER_BOUND]
    assert variable_statuses == [VS.BASIC, VS.BASIC, VS.BASIC, VS.AT_LOWER_BOUND]


def test_compute_statuses_cover_more_cases():  # :2094-2113
    lp = fixtures.make_test_lp()
    lp.variable_upper_bounds = np.array(lp.variable_upper_bounds, dtype=float)
    lp.variable_upper_bounds[3] = 2.5
    lp.variable_upper_bounds[1] = 8
    constraint_statuses, variable_statuses = pdlp_b200.compute_statuses(lp, [-1, 8, 1, 2.5], [-2, 0, 2.375, -1])
    assert constraint_statuses == [CS.FIXED_VALUE, CS.BASIC, CS.AT_LOWER_BOUND, CS.AT_UPPER_BOUND]
    assert variable_statuses == [VS.BASIC, VS.AT_UPPER_BOUND, VS.BASIC, VS.FIXED_VALUE]


def test_compute_statuses_rejects_wrong_sizes():
    with pytest.raises(ValueError):
        pdlp_b200.compute_statuses(fixtures.make_test_lp(), [0, 0, 0], [0, 0, 0, 0])


def test_glop_status_to_termination_reason_table():  # primal_dual_hybrid_gradient.cc:1123-1143
    G, f = pdlp_b200.GlopProblemStatus, pdlp_b200.glop_status_to_termination_reason
    assert f(G.OPTIMAL) == TR.TERMINATION_REASON_OPTIMAL
    assert f(G.INVALID_PROBLEM) == TR.TERMINATION_REASON_INVALID_PROBLEM
    assert f(G.ABNORMAL) == f(G.IMPRECISE) == TR.TERMINATION_REASON_NUMERICAL_ERROR
    for s in (G.PRIMAL_INFEASIBLE, G.DUAL_INFEASIBLE, G.INFEASIBLE_OR_UNBOUNDED, G.DUAL_UNBOUNDED, G.PRIMAL_UNBOUNDED):
        assert f(s) == TR.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE
    for s in (G.INIT, G.PRIMAL_FEASIBLE, G.DUAL_FEASIBLE):
        assert f(s) == TR.TERMINATION_REASON_OTHER
