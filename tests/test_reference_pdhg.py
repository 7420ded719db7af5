"""The reference's end-to-end solver tests (ortools/pdlp/primal_dual_hybrid_gradient_test.cc),
re-typed, run against BOTH drivers of tests/backends.py: the CPU oracle (`-m "not gpu"`,
pins the oracle) and the CUDA path through the C ABI (`-m gpu`).

Not carried over: the Glop presolve cases (:1967-2075, presolve is host-only and outside
the path), DetectsUncompressedConstraintMatrix (:1074, scipy matrices are always
compressed) and the verbosity sweep (:1937, logging only)."""
import itertools
import math

import numpy as np
import pytest
import scipy.sparse as sp

import fixtures
import pdlp_b200

P = pdlp_b200.PrimalDualHybridGradientParams
TR = pdlp_b200.TerminationReason
RC = pdlp_b200.RestartChoice
PT = pdlp_b200.PointType
INF = np.inf

# (enable_scaling, num_threads, use_iteration_limit, malitsky_pock) -- :254-281
LP_COMBOS = list(itertools.product([True, False], [1, 4], [True, False], [True, False]))
LP_IDS = [f"{'Scaling' if s else 'NoScaling'}_{t}T_{'IterLimit' if i else 'NoIterLimit'}_{'MP' if m else 'Adaptive'}" for s, t, i, m in LP_COMBOS]
QP_COMBOS = [c + (d,) for c in LP_COMBOS for d in (True, False)]
QP_IDS = [a + ("_TRDiag" if c[4] else "_TRNoDiag") for a, c in zip([i for i in LP_IDS for _ in (0, 1)], QP_COMBOS)]


def create_solver_params(iteration_limit, eps_abs, combo):
    """CreateSolverParams (:99-135)."""
    scaling, threads, use_limit, mp = combo[:4]
    diag = combo[4] if len(combo) > 4 else False
    p = P()
    if not scaling:
        p.l2_norm_rescaling = False
        p.l_inf_ruiz_iterations = 0
    if mp:
        p.linesearch_rule = P.MALITSKY_POCK_LINESEARCH_RULE
    soc = p.termination_criteria.simple_optimality_criteria
    soc.eps_optimal_relative = 0.0
    if use_limit:
        p.termination_criteria.iteration_limit = iteration_limit
        soc.eps_optimal_absolute = 0.0
    else:
        soc.eps_optimal_absolute = eps_abs
    if diag:
        p.use_diagonal_qp_trust_region_solver = True
        p.diagonal_qp_trust_region_solver_tolerance = 1.0e-8
    p.num_threads = threads
    return p


def verify_termination(params, out, combo, iteration_limit):
    """VerifyTerminationReasonAndIterationCount (:146-177)."""
    log = out.solve_log
    if combo[2]:
        assert log.termination_reason in (TR.TERMINATION_REASON_ITERATION_LIMIT, TR.TERMINATION_REASON_NUMERICAL_ERROR, TR.TERMINATION_REASON_OPTIMAL)
        if log.termination_reason == TR.TERMINATION_REASON_ITERATION_LIMIT:
            assert log.iteration_count == iteration_limit
        else:
            assert log.iteration_count <= iteration_limit
    else:
        # The reference asserts iteration_count <= params.termination_criteria().iteration_limit() here, and
        # CreateSolverParams leaves that field at its default (INT32_MAX) when use_iteration_limit is false: the only
        # requirement of the reference is OPTIMAL.  The second assertion is THIS repository's regression guard: the
        # no-limit solve should not need more iterations than the budget the limit combinations run for.
        assert log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
        assert log.iteration_count <= params.termination_criteria.iteration_limit
        assert log.iteration_count <= iteration_limit


def known_budget_deviation(combo):
    """Where this repository's own regression guard (verify_termination: a no-limit solve should converge within the
    iteration budget its sibling limit combination runs for) is relaxed.  NOT a deviation from the reference: the
    reference asserts no iteration bound at all for the *_NoIterationLimit_* combinations (see verify_termination).

    Without rescaling and with the adaptive step rule (NoScaling_*_NoIterationLimit_AdaptiveLinesearch) the oracle -
    and the device, which follows it - needs more iterations than that budget to reach eps 1e-5 / 1e-8 on two
    fixtures, for EVERY perturbation of the data from 1e-15 to 1e-2 relative (so it is not rounding chaos):
      Tiny (limit combos run 300, :317):                      420 in 72 of 82 perturbed copies, 300 / 360 in the rest;
      InactiveTwoSidedConstraint (limit combos run 500, :395): 600 - 840, never below 540.
    After exactly 300 / 500 iterations the same solves already meet the reference's solution tolerances (the
    *_IterationLimit_* combinations of the same tests pass unchanged), under Malitsky-Pock they stop exactly at 300 and
    480, and with rescaling at 180.  These two combinations carry the oracle's measured envelope as the guard's value."""
    scaling, _, use_limit, mp = combo[:4]
    return (not scaling) and (not use_limit) and (not mp)


def convergence_info(out):
    for c in out.solve_log.solution_stats.convergence_information:
        if c.candidate_type == out.solve_log.solution_type:
            return c
    raise AssertionError("no convergence information for the solution type")


def verify_objectives(out, value, tol):
    c = convergence_info(out)
    assert abs(c.primal_objective - value) <= tol, c.primal_objective
    assert abs(c.dual_objective - value) <= tol, c.dual_objective


def near(a, b, tol):
    np.testing.assert_allclose(np.asarray(a, dtype=float), np.asarray(b, dtype=float), rtol=0, atol=tol)


def check_sizes(out, n, m):
    log = out.solve_log
    assert log.original_problem_stats.num_variables == n and log.preprocessed_problem_stats.num_variables <= n
    assert log.original_problem_stats.num_constraints == m and log.preprocessed_problem_stats.num_constraints <= m


@pytest.mark.parametrize("combo", LP_COMBOS, ids=LP_IDS)
class TestLp:
    def test_unbounded_variables(self, backend, combo):  # :295
        p = create_solver_params(980, 1.0e-7, combo)
        p.major_iteration_frequency = 100
        out = backend.solve(fixtures.make_test_lp(), p)
        verify_termination(p, out, combo, 980)
        verify_objectives(out, -34.0, 1.0e-6)
        near(out.primal_solution, [-1, 8, 1, 2.5], 1.0e-4)
        near(out.dual_solution, [-2, 0, 2.375, 2.0 / 3], 1.0e-4)
        check_sizes(out, 4, 4)

    def test_tiny(self, backend, combo):  # :316
        p = create_solver_params(300, 1.0e-5, combo)
        p.major_iteration_frequency = 60
        budget = 480 if known_budget_deviation(combo) else 300
        out = backend.solve(fixtures.tiny_lp(), p)
        verify_termination(p, out, combo, budget)
        verify_objectives(out, -1.0, 1.0e-4)
        near(out.primal_solution, [1, 0, 6, 2], 1.0e-4)
        near(out.dual_solution, [0.5, 4.0, 0.0], 1.0e-4)
        near(out.reduced_costs, [0.0, 1.5, -3.5, 0.0], 1.0e-4)
        check_sizes(out, 4, 3)

    def test_correlation_clustering_one(self, backend, combo):  # :338
        p = create_solver_params(9, 1.0e-10, combo)
        p.major_iteration_frequency = 2
        out = backend.solve(fixtures.correlation_clustering_lp(), p)
        verify_termination(p, out, combo, 9)
        verify_objectives(out, 1.0, 1.0e-14)
        near(out.primal_solution, [1, 1, 0, 1, 0, 0], 1.0e-14)
        y = out.dual_solution
        assert len(y) == 3 and min(y) >= 0 and y[0] + y[1] >= 1 - 1.0e-14
        assert abs(convergence_info(out).corrected_dual_objective - 1) <= 1.0e-14
        check_sizes(out, 6, 3)

    def test_correlation_clustering_star(self, backend, combo):  # :369
        p = create_solver_params(45, 1.0e-6, combo)
        p.major_iteration_frequency = 5
        out = backend.solve(fixtures.correlation_clustering_star_lp(), p)
        verify_termination(p, out, combo, 45)
        verify_objectives(out, 1.5, 1.0e-6)
        near(out.primal_solution, [0.5, 0.5, 0.5, 0, 0, 0], 1.0e-6)
        near(out.dual_solution, [0.5, 0.5, 0.5], 1.0e-6)
        check_sizes(out, 6, 3)

    def test_inactive_two_sided_constraint(self, backend, combo):  # :394
        p = create_solver_params(500, 1.0e-8, combo)
        p.major_iteration_frequency = 60
        qp = fixtures.make_test_lp()
        qp.constraint_lower_bounds[1] = -10
        budget = 840 if known_budget_deviation(combo) else 500
        out = backend.solve(qp, p)
        verify_termination(p, out, combo, budget)
        verify_objectives(out, -34.0, 1.0e-6)
        near(out.primal_solution, [-1, 8, 1, 2.5], 1.0e-7)
        near(out.dual_solution, [-2, 0, 2.375, 2.0 / 3], 1.0e-7)

    def test_infeasible_primal(self, backend, combo):  # :414
        p = create_solver_params(2000, 1.0e-6, combo)
        p.major_iteration_frequency = 5
        p.termination_criteria.eps_primal_infeasible = 1.0e-6
        out = backend.solve(fixtures.small_primal_infeasible_lp(), p)
        assert out.solve_log.termination_reason == TR.TERMINATION_REASON_PRIMAL_INFEASIBLE
        y = out.dual_solution
        assert abs(y[0] / y[1] - 1) <= 1.0e-6 and y[1] < 0.0
        near(out.reduced_costs, [max(y[1] - y[0], 0.0), max(y[0] - y[1], 0.0)], 1.0e-6)
        assert out.solve_log.iteration_count <= 2000
        check_sizes(out, 2, 2)

    def test_infeasible_dual(self, backend, combo):  # :451
        p = create_solver_params(500, 1.0e-6, combo)
        p.major_iteration_frequency = 5
        out = backend.solve(fixtures.small_dual_infeasible_lp(), p)
        assert out.solve_log.termination_reason == TR.TERMINATION_REASON_DUAL_INFEASIBLE
        x = out.primal_solution
        assert abs(x[0] / x[1] - 1) <= 1.0e-6 and x[1] > 0.0
        assert out.solve_log.iteration_count <= 500

    def test_infeasible_primal_dual(self, backend, combo):  # :471
        p = create_solver_params(600, 1.0e-6, combo)
        p.restart_strategy = P.NO_RESTARTS
        p.major_iteration_frequency = 5
        out = backend.solve(fixtures.small_primal_dual_infeasible_lp(), p)
        assert out.solve_log.termination_reason in (TR.TERMINATION_REASON_DUAL_INFEASIBLE, TR.TERMINATION_REASON_PRIMAL_INFEASIBLE)
        assert out.solve_log.iteration_count <= 600

    def test_lp_without_constraints(self, backend, combo):  # :573
        p = create_solver_params(2, 1.0e-6, combo)
        qp = pdlp_b200.QuadraticProgram()
        qp.resize_and_initialize(3, 0)
        qp.variable_lower_bounds = np.array([-1, -INF, -2.0])
        qp.variable_upper_bounds = np.array([INF, 4, 10.0])
        qp.objective_vector = np.array([1, -1, 2.0])
        out = backend.solve(qp, p)
        verify_termination(p, out, combo, 2)
        verify_objectives(out, -9.0, 1.0e-6)
        near(out.primal_solution, [-1, 4, -2], 1.0e-6)
        assert len(out.dual_solution) == 0
        check_sizes(out, 3, 0)

    def test_lp_without_variables(self, backend, combo):  # :597
        p = create_solver_params(2, 1.0e-6, combo)
        qp = pdlp_b200.QuadraticProgram()
        qp.resize_and_initialize(0, 3)
        qp.constraint_lower_bounds = np.array([-1, -INF, -2.0])
        qp.constraint_upper_bounds = np.array([INF, 4, 10.0])
        out = backend.solve(qp, p)
        verify_termination(p, out, combo, 2)
        verify_objectives(out, 0.0, 1.0e-6)
        assert len(out.primal_solution) == 0
        near(out.dual_solution, [0, 0, 0], 1.0e-6)
        check_sizes(out, 0, 3)

    def test_lp_with_only_fixed_variable(self, backend, combo):  # :618
        p = create_solver_params(2, 0.0, combo)
        qp = pdlp_b200.QuadraticProgram()
        qp.resize_and_initialize(1, 0)
        qp.variable_lower_bounds = np.array([1.0])
        qp.variable_upper_bounds = np.array([1.0])
        qp.objective_vector = np.array([1.0])
        out = backend.solve(qp, p)
        assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
        near(out.primal_solution, [1.0], 1.0e-6)
        assert len(out.dual_solution) == 0 and out.solve_log.iteration_count <= 2

    def test_infeasible_lp_without_variables(self, backend, combo):  # :637
        p = create_solver_params(2, 1.0e-6, combo)
        qp = pdlp_b200.QuadraticProgram()
        qp.resize_and_initialize(0, 1)
        qp.constraint_lower_bounds = np.array([-1.0])
        qp.constraint_upper_bounds = np.array([-1.0])
        out = backend.solve(qp, p)
        assert out.solve_log.termination_reason == TR.TERMINATION_REASON_PRIMAL_INFEASIBLE
        assert len(out.primal_solution) == 0 and out.dual_solution[0] < 0.0
        check_sizes(out, 0, 1)


@pytest.mark.parametrize("combo", QP_COMBOS, ids=QP_IDS)
class TestDiagonalQp:
    def test_diagonal_qp1(self, backend, combo):  # :489
        p = create_solver_params(96, 1.0e-6, combo)
        p.major_iteration_frequency = 12
        out = backend.solve(fixtures.diagonal_qp1(), p)
        verify_termination(p, out, combo, 96)
        verify_objectives(out, 6.0, 1.0e-6)
        near(out.primal_solution, [1.0, 0.0], 1.0e-6)
        near(out.dual_solution, [-1.0], 1.0e-6)
        near(out.reduced_costs, [4.0, 0.0], 1.0e-6)
        check_sizes(out, 2, 1)

    def test_diagonal_qp2(self, backend, combo):  # :510
        p = create_solver_params(240, 1.0e-6, combo)
        p.major_iteration_frequency = 12
        out = backend.solve(fixtures.diagonal_qp2(), p)
        verify_termination(p, out, combo, 240)
        verify_objectives(out, -5.0, 1.0e-6)
        near(out.primal_solution, [3.0, 1.0], 1.0e-6)
        near(out.dual_solution, [0.0], 1.0e-6)
        near(out.reduced_costs, [0.0, 0.0], 1.0e-6)

    def test_diagonal_qp3(self, backend, combo):  # :526
        p = create_solver_params(300, 1.0e-6, combo)
        p.major_iteration_frequency = 15
        out = backend.solve(fixtures.diagonal_qp3(), p)
        verify_termination(p, out, combo, 300)
        verify_objectives(out, 2.0, 1.0e-6)
        near(out.primal_solution, [2.0, 0.0, 1.0], 1.0e-6)
        near(out.dual_solution, [-1.0, 1.0], 1.0e-6)
        near(out.reduced_costs, [0, 0, 0], 1.0e-6)

    def test_qp_warm_start(self, backend, combo):  # :545
        p = create_solver_params(35, 1.0e-6, combo)
        p.major_iteration_frequency = 5
        p.primal_weight_update_smoothing = 0.0
        out = backend.solve(fixtures.diagonal_qp1(), p, initial_solution=([0.999, 0.001], [-0.999]))
        verify_termination(p, out, combo, 35)
        verify_objectives(out, 6.0, 1.0e-6)
        near(out.primal_solution, [1.0, 0.0], 1.0e-6)
        near(out.dual_solution, [-1.0], 1.0e-6)
        near(out.reduced_costs, [4.0, 0.0], 1.0e-6)
