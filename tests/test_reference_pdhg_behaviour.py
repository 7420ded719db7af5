"""The un-parametrised solver tests of ortools/pdlp/primal_dual_hybrid_gradient_test.cc
(:672-1935): restart bookkeeping, logging, validation, warm starts, callbacks, interrupts and
feasibility polishing.  Same two drivers as test_reference_pdhg.py."""
import math

import numpy as np
import pytest
import scipy.sparse as sp

import fixtures
import pdlp_b200
from test_reference_pdhg import convergence_info, create_solver_params, near, verify_objectives, verify_termination

P = pdlp_b200.PrimalDualHybridGradientParams
TR = pdlp_b200.TerminationReason
RC = pdlp_b200.RestartChoice
PT = pdlp_b200.PointType
INF = np.inf
NAN = math.nan


def params_with_no_limits():  # :658-670
    p = P()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    p.record_iteration_stats = True
    return p


def by_type(items, t, field="candidate_type"):
    return [c for c in items if getattr(c, field) == t]


@pytest.mark.parametrize("strategy,expected", [(P.NO_RESTARTS, RC.RESTART_CHOICE_WEIGHTED_AVERAGE_RESET), (P.EVERY_MAJOR_ITERATION, RC.RESTART_CHOICE_RESTART_TO_AVERAGE)])
def test_restart_bookkeeping_every_major_iteration(backend, strategy, expected):  # :672-733
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 100
    p.termination_check_frequency = 1
    p.major_iteration_frequency = 17
    p.restart_strategy = strategy
    out = backend.solve(fixtures.make_test_lp(), p)
    assert out.solve_log.iteration_count == 100 and len(out.solve_log.iteration_stats) == 101
    for i, stats in enumerate(out.solve_log.iteration_stats):
        assert stats.iteration_number == i
        want = expected if (i > 0 and i % 17 == 0) else RC.RESTART_CHOICE_NO_RESTART
        assert stats.restart_used == want, i


def test_solve_log_name_and_params(backend):  # :735-763
    p = P()
    p.termination_criteria.iteration_limit = 1
    qp = fixtures.make_test_lp()
    qp.problem_name = "Test LP"
    out = backend.solve(qp, p)
    assert out.solve_log.instance_name == "Test LP"
    out = backend.solve(fixtures.make_test_lp(), p)
    assert not out.solve_log.HasField("instance_name")
    assert out.solve_log.params.termination_criteria.iteration_limit == 1


def test_adaptive_distance_based_restarts(backend):  # :765-798
    p = P()
    p.major_iteration_frequency = 16
    p.termination_criteria.iteration_limit = 128
    p.restart_strategy = P.ADAPTIVE_DISTANCE_BASED
    p.necessary_reduction_for_restart = 0.99
    out = backend.solve(fixtures.make_test_lp(), p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    near(out.primal_solution, [-1, 8, 1, 2.5], 1.0e-4)
    near(out.dual_solution, [-2, 0, 2.375, 2.0 / 3], 1.0e-4)
    verify_objectives(out, -34.0, 1.0e-4)
    out = backend.solve(fixtures.diagonal_qp1(), p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def test_adaptive_distance_based_restarts_to_average(backend):  # :800-835
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 100
    p.termination_check_frequency = 1
    p.major_iteration_frequency = 13
    p.restart_strategy = P.ADAPTIVE_DISTANCE_BASED
    p.necessary_reduction_for_restart = 0.75
    out = backend.solve(fixtures.make_test_lp(), p)
    assert out.solve_log.iteration_count == 100 and len(out.solve_log.iteration_stats) == 101
    for i, stats in enumerate(out.solve_log.iteration_stats):
        assert stats.iteration_number == i
        if i == 0:
            assert stats.restart_used == RC.RESTART_CHOICE_NO_RESTART
        elif i == 13:
            assert stats.restart_used in (RC.RESTART_CHOICE_RESTART_TO_AVERAGE, RC.RESTART_CHOICE_WEIGHTED_AVERAGE_RESET)
        elif i % 13 != 0:
            assert stats.restart_used == RC.RESTART_CHOICE_NO_RESTART


def test_primal_weight_frozen(backend):  # :837-856
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 100
    p.major_iteration_frequency = 17
    p.restart_strategy = P.EVERY_MAJOR_ITERATION
    p.initial_primal_weight = 1.5
    p.primal_weight_update_smoothing = 0.0
    out = backend.solve(fixtures.make_test_lp(), p)
    assert len(out.solve_log.iteration_stats) > 0
    for stats in out.solve_log.iteration_stats:
        assert stats.primal_weight == 1.5, stats.iteration_number


def test_constant_step_size_and_scaling(backend):  # :858-899
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 100
    p.termination_check_frequency = 1
    p.linesearch_rule = P.CONSTANT_STEP_SIZE_RULE
    out = backend.solve(fixtures.make_test_lp(), p)
    steps = [s.step_size for s in out.solve_log.iteration_stats]
    assert len(steps) == 101 and all(s == steps[0] for s in steps)
    p.termination_criteria.iteration_limit = 1
    base = backend.solve(fixtures.make_test_lp(), p).solve_log.iteration_stats[0].step_size
    p.initial_step_size_scaling = 0.5
    assert backend.solve(fixtures.make_test_lp(), p).solve_log.iteration_stats[0].step_size == base * 0.5


def test_kkt_matrix_pass_termination(backend):  # :901-916
    p = params_with_no_limits()
    p.termination_criteria.kkt_matrix_pass_limit = 13
    p.linesearch_rule = P.CONSTANT_STEP_SIZE_RULE
    out = backend.solve(fixtures.make_test_lp(), p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT
    assert out.solve_log.solution_stats.cumulative_kkt_matrix_passes == 13


def test_stats_at_each_iteration(backend):  # :918-955
    p = params_with_no_limits()
    p.termination_check_frequency = 1
    p.termination_criteria.iteration_limit = 100
    p.major_iteration_frequency = 17
    out = backend.solve(fixtures.make_test_lp(), p)
    assert len(out.solve_log.iteration_stats) == 101
    for stats in out.solve_log.iteration_stats:
        assert by_type(stats.convergence_information, PT.POINT_TYPE_CURRENT_ITERATE)
        assert by_type(stats.infeasibility_information, PT.POINT_TYPE_CURRENT_ITERATE)
        assert by_type(stats.point_metadata, PT.POINT_TYPE_CURRENT_ITERATE, "point_type")
        if stats.iteration_number > 0:
            assert by_type(stats.convergence_information, PT.POINT_TYPE_AVERAGE_ITERATE)
            assert by_type(stats.infeasibility_information, PT.POINT_TYPE_AVERAGE_ITERATE)
            assert by_type(stats.point_metadata, PT.POINT_TYPE_AVERAGE_ITERATE, "point_type")
            assert by_type(stats.infeasibility_information, PT.POINT_TYPE_ITERATE_DIFFERENCE)
            assert by_type(stats.point_metadata, PT.POINT_TYPE_ITERATE_DIFFERENCE, "point_type")


def test_no_iteration_stats_when_off(backend):  # :957-972
    p = params_with_no_limits()
    p.record_iteration_stats = False
    p.termination_criteria.iteration_limit = 100
    p.major_iteration_frequency = 17
    p.random_projection_seeds.append(1)
    assert len(backend.solve(fixtures.make_test_lp(), p).solve_log.iteration_stats) == 0


@pytest.mark.parametrize("seeds", [[], [1, 2]])
def test_random_projections(backend, seeds):  # :974-1017
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 100
    p.major_iteration_frequency = 17
    p.termination_check_frequency = 1
    p.random_projection_seeds.extend(seeds)
    out = backend.solve(fixtures.make_test_lp(), p)
    assert len(out.solve_log.iteration_stats) == 101
    for stats in out.solve_log.iteration_stats:
        assert len(stats.point_metadata) > 0
        for md in stats.point_metadata:
            assert len(md.random_primal_projections) == len(seeds) and len(md.random_dual_projections) == len(seeds)


def test_project_initial_point(backend):  # :1019-1047
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 5
    out = backend.solve(fixtures.small_initialization_lp(), p)
    assert len(out.solve_log.iteration_stats) == 6 and out.primal_solution[0] > 0.0
    out = backend.solve(fixtures.small_initialization_lp(), p, initial_solution=([1.0, 0.0], [-1.0, -1.0]))
    assert len(out.solve_log.iteration_stats) == 6 and out.dual_solution[0] <= 0.0


def _test_lp_with(**changes):
    qp = fixtures.make_test_lp()
    for name, (index, value) in changes.items():
        getattr(qp, name)[index] = value
    return qp


def _with_matrix_entry(row, col, value):
    qp = fixtures.make_test_lp()
    K = qp.constraint_matrix.tolil()
    K[row, col] = value
    qp.constraint_matrix = K.tocsc()
    return qp


def _qp1_with_diag0(value):
    qp = fixtures.diagonal_qp1()
    d = np.array(qp.objective_matrix, dtype=float)
    d[0] = value
    qp.set_objective_matrix_diagonal(d)
    return qp


def _zero_scaling():
    qp = fixtures.make_test_lp()
    qp.objective_scaling_factor = 0.0
    return qp


def _inconsistent_sizes():
    qp = fixtures.tiny_lp()
    qp.objective_vector = np.zeros(0)
    return qp


INVALID_PROBLEMS = {  # :1049-1291,1372-1380
    "inconsistent_bounds": fixtures.small_invalid_problem_lp,
    "nonconvex_qp": lambda: _qp1_with_diag0(-1.0),
    "inconsistent_sizes": _inconsistent_sizes,
    "nan_in_matrix": lambda: _with_matrix_entry(0, 0, NAN),
    "excessive_matrix": lambda: _with_matrix_entry(0, 0, 1.0e60),
    "tiny_col_norm": lambda: _with_matrix_entry(0, 1, 1.0e-60),
    "tiny_row_norm": lambda: _with_matrix_entry(2, 0, 1.0e-60),
    "nan_constraint_bound": lambda: _test_lp_with(constraint_upper_bounds=(1, NAN)),
    "excessive_constraint_ub": lambda: _test_lp_with(constraint_upper_bounds=(1, 1.0e60)),
    "excessive_constraint_lb": lambda: _test_lp_with(constraint_lower_bounds=(2, -1.0e60)),
    "nan_variable_bound": lambda: _test_lp_with(variable_lower_bounds=(3, NAN)),
    "excessive_variable_bound": lambda: _test_lp_with(variable_lower_bounds=(3, -1.0e60)),
    "nan_objective": lambda: _test_lp_with(objective_vector=(3, NAN)),
    "excessive_objective": lambda: _test_lp_with(objective_vector=(3, -1.0e60)),
    "nan_objective_matrix": lambda: _qp1_with_diag0(NAN),
    "excessive_objective_matrix": lambda: _qp1_with_diag0(1.0e60),
    "zero_objective_scaling_factor": _zero_scaling,
}


@pytest.mark.parametrize("case", sorted(INVALID_PROBLEMS))
def test_detects_invalid_problem(backend, case):
    out = backend.solve(INVALID_PROBLEMS[case](), P())
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PROBLEM, out.solve_log.termination_string


def _tiny_ub():
    qp = fixtures.make_test_lp()
    qp.variable_lower_bounds[1] = 0.0
    qp.variable_upper_bounds[1] = 1.0e-60
    return qp


SMALL_VALUES_OK = {  # :1148-1262 (without presolve)
    "constraint_ub": lambda: _test_lp_with(constraint_upper_bounds=(1, 1.0e-60)),
    "constraint_lb": lambda: _test_lp_with(constraint_lower_bounds=(2, 1.0e-60)),
    "variable_bound": _tiny_ub,
    "objective": lambda: _test_lp_with(objective_vector=(3, 1.0e-60)),
}


@pytest.mark.parametrize("case", sorted(SMALL_VALUES_OK))
def test_excessively_small_values_ok_without_presolve(backend, case):
    out = backend.solve(SMALL_VALUES_OK[case](), P())
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def test_detects_invalid_parameters(backend):  # :1083-1089,1382-1390
    p = P()
    p.num_threads = 0
    assert backend.solve(fixtures.tiny_lp(), p).solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PARAMETER
    p = P()
    p.use_feasibility_polishing = True
    assert backend.solve(fixtures.diagonal_qp1(), p).solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PARAMETER


@pytest.mark.parametrize("x0,y0", [
    ([1.0, NAN, 1.0, 1.0], [1.0, 1.0, 1.0, 1.0]), ([1.0, 1.0e100, 1.0, 1.0], [1.0, 1.0, 1.0, 1.0]), ([1.0, 1.0, 1.0], [1.0, 1.0, 1.0, 1.0]),
    ([1.0, 1.0, 1.0, 1.0], [1.0, NAN, 1.0, 1.0]), ([1.0, 1.0, 1.0, 1.0], [1.0, 1.0e100, 1.0, 1.0]), ([1.0, 1.0, 1.0, 1.0], [1.0, 1.0, 1.0]),
])
def test_detects_invalid_initial_solution(backend, x0, y0):  # :1293-1370
    out = backend.solve(fixtures.make_test_lp(), P(), initial_solution=(x0, y0))
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_INITIAL_SOLUTION


def test_checks_termination_at_correct_frequency(backend):  # :1392-1419
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 16
    p.termination_check_frequency = 2
    p.major_iteration_frequency = 5
    out = backend.solve(fixtures.make_test_lp(), p)
    assert out.solve_log.iteration_count == 16
    checked = [s.iteration_number for s in out.solve_log.iteration_stats if len(s.convergence_information) > 0]
    assert checked == [0, 2, 4, 5, 7, 9, 10, 12, 14, 15, 16]


def test_calls_callback(backend):  # :1421-1443
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 16
    p.termination_check_frequency = 5
    p.major_iteration_frequency = 5
    calls = []
    out = backend.solve(fixtures.make_test_lp(), p, callback=lambda t, stats: calls.append((t, stats.iteration_number)))
    assert out.solve_log.iteration_count == 16
    assert len(calls) == 6  # iterations 0, 5, 10, 15, 16 and the final solution
    assert [n for _, n in calls[:5]] == [0, 5, 10, 15, 16]


def test_warm_started_at_optimum(backend):  # :1453-1469
    p = params_with_no_limits()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1.0e-10
    out = backend.solve(fixtures.tiny_lp(), p, initial_solution=([1.0, 0.0, 6.0, 2.0], [0.5, 4.0, 0.0]))
    near(out.primal_solution, [1, 0, 6, 2], 1.0e-10)
    near(out.dual_solution, [0.5, 4.0, 0.0], 1.0e-10)
    assert out.solve_log.iteration_count <= 0 and out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def test_no_movement_when_started_at_optimum(backend):  # :1471-1485
    p = params_with_no_limits()
    p.l2_norm_rescaling = False
    p.l_inf_ruiz_iterations = 0
    out = backend.solve(fixtures.tiny_lp(), p, initial_solution=([1.0, 0.0, 6.0, 2.0], [0.5, 4.0, 0.0]))
    assert list(out.primal_solution) == [1, 0, 6, 2] and list(out.dual_solution) == [0.5, 4.0, 0.0]
    assert out.solve_log.iteration_count <= 1 and out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def test_empty_qp(backend):  # :1487-1498
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(0, 0)
    out = backend.solve(qp, params_with_no_limits())
    assert len(out.primal_solution) == 0 and len(out.dual_solution) == 0
    assert out.solve_log.iteration_count == 0 and out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def _zero_eps():
    p = P()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    return p


def test_respects_interrupt(backend):  # :1500-1515
    flag = np.ones(1, dtype=np.int32)
    out = backend.solve(fixtures.make_test_lp(), _zero_eps(), interrupt=flag)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INTERRUPTED_BY_USER


def test_respects_interrupt_from_callback(backend):  # :1517-1539
    flag = np.zeros(1, dtype=np.int32)

    def callback(_type, stats):
        if stats.iteration_number >= 10:
            flag[0] = 1

    out = backend.solve(fixtures.make_test_lp(), _zero_eps(), interrupt=flag, callback=callback)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INTERRUPTED_BY_USER
    assert out.solve_log.iteration_count >= 10


def test_ignores_false_interrupt(backend):  # :1541-1557
    p = _zero_eps()
    p.termination_criteria.kkt_matrix_pass_limit = 1
    flag = np.zeros(1, dtype=np.int32)
    out = backend.solve(fixtures.make_test_lp(), p, interrupt=flag)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT


@pytest.mark.parametrize("field", ["num_threads", "num_shards"])
def test_huge_num_threads_or_shards(backend, field):  # :1559-1581
    p = params_with_no_limits()
    p.termination_criteria.iteration_limit = 10
    setattr(p, field, 1_000_000_000)
    assert backend.solve(fixtures.make_test_lp(), p).solve_log.termination_reason == TR.TERMINATION_REASON_ITERATION_LIMIT


def test_detailed_termination_criteria(backend):  # :1583-1613
    combo = (True, 4, False, False)
    p = create_solver_params(300, 1.0e-5, combo)
    p.major_iteration_frequency = 60
    p.termination_criteria.ClearField("simple_optimality_criteria")
    d = p.termination_criteria.detailed_optimality_criteria
    d.eps_optimal_primal_residual_absolute = 1.0e-5
    d.eps_optimal_primal_residual_relative = 0.0
    d.eps_optimal_dual_residual_absolute = 1.0e-5
    d.eps_optimal_dual_residual_relative = 0.0
    d.eps_optimal_objective_gap_absolute = 1.0e-5
    d.eps_optimal_objective_gap_relative = 0.0
    out = backend.solve(fixtures.tiny_lp(), p)
    verify_termination(p, out, combo, 300)
    verify_objectives(out, -1.0, 1.0e-4)
    near(out.primal_solution, [1, 0, 6, 2], 1.0e-4)
    near(out.dual_solution, [0.5, 4.0, 0.0], 1.0e-4)
    near(out.reduced_costs, [0.0, 1.5, -3.5, 0.0], 1.0e-4)


# ---- feasibility polishing (:1615-1935) ----
def polishing_params():
    p = P()
    p.linesearch_rule = P.CONSTANT_STEP_SIZE_RULE
    d = p.termination_criteria.detailed_optimality_criteria
    d.eps_optimal_primal_residual_absolute = 1.0e-6
    d.eps_optimal_primal_residual_relative = 1.0e-6
    d.eps_optimal_dual_residual_absolute = 1.0e-6
    d.eps_optimal_dual_residual_relative = 1.0e-6
    d.eps_optimal_objective_gap_absolute = 1.0e-2
    d.eps_optimal_objective_gap_relative = 1.0e-2
    p.termination_criteria.iteration_limit = 500
    p.handle_some_primal_gradients_on_finite_bounds_as_residuals = False
    p.primal_weight_update_smoothing = 0.0
    p.use_feasibility_polishing = True
    return p


def polishing_primal_lp():
    # min x_1 + 1.001 x_2 s.t. x_1 + x_2 = 1, x >= -1
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(2, 1)
    qp.objective_vector = np.array([1.0, 1.001])
    qp.variable_lower_bounds = np.array([-1.0, -1.0])
    qp.variable_upper_bounds = np.array([INF, INF])
    qp.constraint_lower_bounds = np.array([1.0])
    qp.constraint_upper_bounds = np.array([1.0])
    qp.constraint_matrix = sp.csc_matrix(np.array([[1.0, 1.0]]))
    return qp


def polishing_dual_lp():
    # min -y s.t. y <= 1, y <= 1.001
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(1, 2)
    qp.objective_vector = np.array([-1.0])
    qp.constraint_lower_bounds = np.array([-INF, -INF])
    qp.constraint_upper_bounds = np.array([1.0, 1.001])
    qp.constraint_matrix = sp.csc_matrix(np.array([[1.0], [1.0]]))
    return qp


PH_PRIMAL, PH_DUAL = 1, 2  # POLISHING_PHASE_TYPE_* (solve_log.proto)


@pytest.mark.parametrize("make", [polishing_primal_lp, polishing_dual_lp], ids=["primal", "dual"])
def test_feasibility_polishing_solves_faster(backend, make):
    p = polishing_params()
    p.use_feasibility_polishing = False
    assert backend.solve(make(), p).solve_log.termination_reason == TR.TERMINATION_REASON_ITERATION_LIMIT
    p.use_feasibility_polishing = True
    assert backend.solve(make(), p).solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL


def test_feasibility_polishing_finds_valid_solution(backend):
    out = backend.solve(polishing_primal_lp(), polishing_params())
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    verify_objectives(out, 1.0, 1.0e-2)
    x, y = out.primal_solution, out.dual_solution
    assert abs(x[0] + x[1] - 1.0) <= 1.0e-6 and x[0] >= 0.0 and x[1] >= 0.0
    assert len(y) == 1 and y[0] <= 1.0 and y[0] >= 1.0 - 1e-2
    near(out.reduced_costs, [1.0 - y[0], 1.001 - y[0]], 1.0e-12)


def test_feasibility_polishing_details_in_log(backend):
    """FeasibilityPolishingDetailsInLog (:1740), ...UsesInfiniteTolerances (:1775), ...WorkStatsAreCorrect (:1816) and
    FeasibilityObjectivesAreZero (:1857) on one solve."""
    p = polishing_params()
    p.record_iteration_stats = True
    out = backend.solve(polishing_primal_lp(), p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    primal_optimal = dual_optimal = total_iterations = 0
    primal_infos = dual_infos = 0
    total_feasibility_time = 0.0
    for phase in out.solve_log.feasibility_polishing_details:
        total_feasibility_time += phase.solve_time_sec
        assert phase.polishing_phase_type in (PH_PRIMAL, PH_DUAL)
        assert phase.HasField("solution_stats") and phase.solution_stats.iteration_number == phase.iteration_count
        assert 0 <= phase.iteration_count <= phase.main_iteration_count / 4
        total_iterations += phase.iteration_count
        crit = phase.params.termination_criteria
        assert crit.HasField("detailed_optimality_criteria")
        d = crit.detailed_optimality_criteria
        assert d.eps_optimal_objective_gap_absolute == INF and d.eps_optimal_objective_gap_relative == INF
        if phase.polishing_phase_type == PH_PRIMAL:
            primal_optimal += phase.termination_reason == TR.TERMINATION_REASON_OPTIMAL
            assert d.eps_optimal_dual_residual_absolute == INF and d.eps_optimal_dual_residual_relative == INF
        else:
            dual_optimal += phase.termination_reason == TR.TERMINATION_REASON_OPTIMAL
            assert d.eps_optimal_primal_residual_absolute == INF and d.eps_optimal_primal_residual_relative == INF
        for it in phase.iteration_stats:
            for c in it.convergence_information:
                if phase.polishing_phase_type == PH_PRIMAL:
                    primal_infos += 1
                    assert c.primal_objective == 0.0
                else:
                    dual_infos += 1
                    assert c.dual_objective == 0.0
    assert primal_optimal >= 1 and dual_optimal >= 1 and primal_infos >= 1 and dual_infos >= 1
    last = out.solve_log.iteration_stats[-1]
    assert total_iterations >= 1 and last.iteration_number >= 1
    assert last.iteration_number + total_iterations == out.solve_log.iteration_count
    # solve_time_sec covers the main and the feasibility iterations; the main loop's cumulative_time_sec does not
    # include the phases' time (:1843-1853).
    assert total_feasibility_time > 0.0 and last.cumulative_time_sec > 0.0
    assert total_feasibility_time + last.cumulative_time_sec <= out.solve_log.solve_time_sec


def test_feasibility_polishing_callbacks(backend):
    counts = {}
    order = []

    def callback(t, _stats):
        counts[t] = counts.get(t, 0) + 1
        order.append(t)

    backend.solve(polishing_primal_lp(), polishing_params(), callback=callback)
    # IterationType: kNormal 0, kPrimalFeasibility 1, kDualFeasibility 2, kPresolveTermination 3,
    # kNormalTermination 4, kFeasibilityPolishingTermination 5 (primal_dual_hybrid_gradient.h:70-83)
    assert order[-1] == 5
    assert sorted(counts) == [0, 1, 2, 5] and all(v >= 1 for v in counts.values())


def test_iteration_limit_includes_feasibility_phases(backend):
    p = polishing_params()
    p.termination_criteria.iteration_limit = 300
    p.record_iteration_stats = True
    out = backend.solve(fixtures.tiny_lp(), p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    polishing_iterations = sum(ph.iteration_count for ph in out.solve_log.feasibility_polishing_details)
    assert polishing_iterations >= 12
    assert out.solve_log.iteration_stats[-1].iteration_number + polishing_iterations == out.solve_log.iteration_count
