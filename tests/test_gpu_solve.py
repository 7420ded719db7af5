"""End-to-end parity on the GPU: pdlp_b200 (CUDA path through the C ABI) against the CPU
oracle on the reference's fixtures (primal_dual_hybrid_gradient_test.cc:295-541) and on
seeded synthetic LPs.  Tolerances from BASELINE.json north_star: same termination
reason, objectives within 1e-6 relative, iteration count within 5 %."""
import numpy as np
import pytest

import fixtures
import oracle_binding as ob
import pdlp_b200

pytestmark = pytest.mark.gpu

P = pdlp_b200.PrimalDualHybridGradientParams


def params(eps_abs=1e-7, eps_rel=0.0, **kw):
    p = P()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = eps_abs
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = eps_rel
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def both(qp, p, **kw):
    dev = pdlp_b200.primal_dual_hybrid_gradient(qp, p, **kw)
    init = kw.get("initial_solution")
    ora = ob.solve(qp, p, None if init is None else (init.primal_solution, init.dual_solution))
    return dev, ora


def assert_parity(dev, ora, obj_tol=1e-6, iter_tol=0.05, iter_slack=2):
    assert dev.solve_log.termination_reason == ora.solve_log_dict["termination_reason"]
    di, oi = dev.solve_log.iteration_count, ora.solve_log_dict["iteration_count"]
    assert abs(di - oi) <= max(iter_slack, iter_tol * oi), (di, oi)
    if ora.solve_log_dict["solution_stats"]["convergence_information"]:
        oc = ora.solve_log_dict["solution_stats"]["convergence_information"][0]
        dc = dev.solve_log.solution_stats.convergence_information[0]
        for name in ("primal_objective", "dual_objective"):
            a, b = getattr(dc, name), oc[name]
            assert abs(a - b) <= obj_tol * max(1.0, abs(b)), (name, a, b)


FIXTURES = [
    ("test_lp", fixtures.make_test_lp, -34.0, [-1, 8, 1, 2.5], [-2, 0, 2.375, 2.0 / 3], 980, 1e-7),
    ("tiny_lp", fixtures.tiny_lp, -1.0, [1, 0, 6, 2], [0.5, 4, 0], 300, 1e-5),
    ("cc_lp", fixtures.correlation_clustering_lp, 1.0, [1, 1, 0, 1, 0, 0], None, 9, 1e-10),
    ("cc_star", fixtures.correlation_clustering_star_lp, 1.5, [0.5, 0.5, 0.5, 0, 0, 0], [0.5, 0.5, 0.5], 45, 1e-6),
    ("qp1", fixtures.diagonal_qp1, 6.0, [1, 0], [-1], 96, 1e-6),
    ("qp2", fixtures.diagonal_qp2, -5.0, [3, 1], [0], 240, 1e-6),
    ("qp3", fixtures.diagonal_qp3, 2.0, [2, 0, 1], [-1, 1], 300, 1e-6),
]


@pytest.mark.parametrize("name,make,objective,x,y,budget,eps", FIXTURES, ids=[f[0] for f in FIXTURES])
def test_known_answer_fixtures(name, make, objective, x, y, budget, eps):
    qp = make()
    # major_iteration_frequency as the reference's own tests set it (primal_dual_hybrid_gradient_test.cc:300-530)
    extra = {"major_iteration_frequency": {"test_lp": 100, "tiny_lp": 60, "cc_lp": 2, "cc_star": 5, "qp1": 12, "qp2": 12, "qp3": 15}[name]}
    dev, ora = both(qp, params(eps_abs=eps, **extra))
    assert dev.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    assert dev.solve_log.iteration_count <= budget  # the reference's own iteration budgets
    tol = max(1e-5, 100 * eps)
    np.testing.assert_allclose(dev.primal_solution, x, atol=tol)
    if y is not None:
        np.testing.assert_allclose(dev.dual_solution, y, atol=tol)
        np.testing.assert_allclose(dev.dual_solution, ora.dual_solution, atol=tol)
    ci = dev.solve_log.solution_stats.convergence_information[0]
    assert abs(ci.primal_objective - objective) < tol
    assert_parity(dev, ora)
    np.testing.assert_allclose(dev.primal_solution, ora.primal_solution, atol=tol)
    np.testing.assert_allclose(dev.reduced_costs, ora.reduced_costs, atol=10 * tol)


@pytest.mark.parametrize("rule", [P.ADAPTIVE_LINESEARCH_RULE, P.MALITSKY_POCK_LINESEARCH_RULE, P.CONSTANT_STEP_SIZE_RULE])
@pytest.mark.parametrize("scaling", [True, False])
def test_step_rules_and_scaling_on_test_lp(rule, scaling):
    p = params(eps_abs=1e-6, linesearch_rule=rule)
    if not scaling:
        p.l_inf_ruiz_iterations = 0
        p.l2_norm_rescaling = False
    dev, ora = both(fixtures.make_test_lp(), p)
    assert dev.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    np.testing.assert_allclose(dev.primal_solution, [-1, 8, 1, 2.5], atol=1e-4)
    if rule != P.CONSTANT_STEP_SIZE_RULE:  # the constant rule's sigma_max start vector is parity-unpinned
        assert_parity(dev, ora)


def assert_objectives_close(dev, ora, tol=1e-6):
    oc = ora.solve_log_dict["solution_stats"]["convergence_information"][0]
    dc = dev.solve_log.solution_stats.convergence_information[0]
    for name in ("primal_objective", "dual_objective"):
        a, b = getattr(dc, name), oc[name]
        assert abs(a - b) <= tol * max(1.0, abs(b)), (name, a, b)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_lp_parity(seed):
    qp = fixtures.random_lp(400, 800, 5, seed)
    p = params(eps_abs=1e-7, eps_rel=1e-7)  # both solvers then pin the objective to well under 1e-6 relative
    dev, ora = both(qp, p)
    assert dev.solve_log.termination_reason == ora.solve_log_dict["termination_reason"] == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    assert_objectives_close(dev, ora)  # iteration counts: tests/test_gpu_iteration_parity.py (a distribution, not one pair of solves)
    ci = [c for c in dev.solve_log.solution_stats.convergence_information if c.candidate_type == dev.solve_log.solution_type][0]
    bn = np.linalg.norm(np.where(np.isfinite(qp.constraint_upper_bounds), qp.constraint_upper_bounds, 0.0))
    assert ci.l2_primal_residual <= 1e-7 * (1 + bn) * 1.0001


def test_random_qp_parity():
    qp = fixtures.random_lp(300, 600, 5, 3, qp_diagonal=True)
    p = params(eps_abs=1e-7, eps_rel=1e-7)
    dev, ora = both(qp, p)
    assert dev.solve_log.termination_reason == ora.solve_log_dict["termination_reason"] == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    assert_objectives_close(dev, ora)


def test_trace_on_random_lp_follows_the_oracle_until_rounding_takes_over():
    """On a 400 x 800 LP the device and the oracle start from identical arithmetic; restart tags
    and rejected-step counts must agree and step sizes stay within 1e-6 for at least the first
    64 iterations (rounding differences grow ~1e3 x per 50 iterations after that)."""
    qp = fixtures.random_lp(400, 800, 5, 1)
    p = params(eps_abs=0.0, record_iteration_stats=True)
    p.termination_criteria.iteration_limit = 256
    dev, ora = both(qp, p)
    ds, os_ = dev.solve_log.iteration_stats, ora.solve_log.iteration_stats
    assert len(ds) == len(os_) == 257
    agree = 0
    for d, o in zip(ds, os_):
        if d.restart_used != o.restart_used or d.cumulative_rejected_steps != o.cumulative_rejected_steps:
            break
        if abs(d.step_size - o.step_size) > 1e-6 * o.step_size or abs(d.primal_weight - o.primal_weight) > 1e-6 * o.primal_weight:
            break
        agree += 1
    print("device trace follows the oracle for", agree, "iterations")
    assert agree >= 64, agree


def test_iteration_limit_trace_matches_oracle():
    """With record_iteration_stats every iteration's step size / primal weight / restart
    tag is compared against the oracle."""
    p = params(eps_abs=0.0, record_iteration_stats=True)
    p.termination_criteria.iteration_limit = 60
    dev, ora = both(fixtures.make_test_lp(), p)
    assert dev.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_ITERATION_LIMIT
    ds, os_ = dev.solve_log.iteration_stats, ora.solve_log_dict["iteration_stats"]
    assert len(ds) == len(os_)
    for d, o in zip(ds, os_):
        assert d.iteration_number == o["iteration_number"]
        assert d.restart_used == o["restart_used"], d.iteration_number
        assert d.cumulative_rejected_steps == o["cumulative_rejected_steps"]
        np.testing.assert_allclose(d.step_size, o["step_size"], rtol=1e-9)
        np.testing.assert_allclose(d.primal_weight, o["primal_weight"], rtol=1e-9)
