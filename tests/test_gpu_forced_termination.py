"""Forced numerical termination on the device against the oracle (`-m gpu`): a trial step with zero movement (a fixed
point reached between two termination checks, primal_dual_hybrid_gradient.cc:2313-2317) and one whose movement
overflows (:2318-2322).  The reference's `for (iterations_completed_ = 0;; ++iterations_completed_)` (:2716) counts the
iteration whose step forced the termination before the final check sees it; iteration_count, rejected steps and KKT
passes of the result must agree with the oracle's."""
import numpy as np
import pytest
import scipy.sparse as sp

import oracle_binding as ob
import pdlp_b200

pytestmark = pytest.mark.gpu

TR = pdlp_b200.TerminationReason
INF = np.inf


def _qp(K, c, vlb, vub, clb, cub):
    q = pdlp_b200.QuadraticProgram()
    q.constraint_matrix = sp.csc_matrix(np.array(K, dtype=float))
    q.objective_vector = np.array(c, dtype=float)
    q.variable_lower_bounds, q.variable_upper_bounds = np.array(vlb, dtype=float), np.array(vub, dtype=float)
    q.constraint_lower_bounds, q.constraint_upper_bounds = np.array(clb, dtype=float), np.array(cub, dtype=float)
    return q


def _params(**kw):
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    p.l_inf_ruiz_iterations = 0
    p.l2_norm_rescaling = False
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _both(qp, p, init=None):
    ora = ob.solve(qp, p, init).solve_log_dict
    start = None if init is None else pdlp_b200.PrimalAndDualSolution(*init)
    dev = pdlp_b200.primal_dual_hybrid_gradient(qp, p, initial_solution=start, device=0).solve_log
    return dev, ora


def _assert_same_work(dev, ora):
    assert dev.termination_reason == ora["termination_reason"]
    assert dev.iteration_count == ora["iteration_count"]
    s = ora["solution_stats"]
    assert dev.solution_stats.cumulative_rejected_steps == s["cumulative_rejected_steps"]
    assert dev.solution_stats.cumulative_kkt_matrix_passes == s["cumulative_kkt_matrix_passes"]


def test_fixed_point_between_two_checks_counts_its_iteration():
    # min x, x in [0, 1], started at 0.5: the first step lands on the bound, the second does not move.
    qp = _qp([[1.0]], [1.0], [0.0], [1.0], [-INF], [INF])
    dev, ora = _both(qp, _params(), (np.array([0.5]), np.array([0.0])))
    assert ora["termination_reason"] == TR.TERMINATION_REASON_OPTIMAL and ora["iteration_count"] == 2
    _assert_same_work(dev, ora)


def test_overflowing_movement_is_a_numerical_error_after_one_iteration():
    qp = _qp([[1e30, 1e30], [1e30, -1e30]], [9e49, -9e49], [-INF, -INF], [INF, INF], [-1e30, -1e30], [1e30, 1e30])
    dev, ora = _both(qp, _params(initial_step_size_scaling=1e40))
    assert ora["termination_reason"] == TR.TERMINATION_REASON_NUMERICAL_ERROR and ora["iteration_count"] == 1
    _assert_same_work(dev, ora)
