"""Cases added after this round's GPU budget was spent: their oracle variants run in the CPU suite, and their device
variants run there too on the whole-library emulation (tests/test_full_emulation.py: the host driver and the kernels as
host C++) - but not yet on a B200.  The file sorts after every test file that has run on the hardware (and before the
allocation test, which must be last in its process) so that a surprise there cannot hide results that were green."""
import numpy as np
import pytest

import pdlp_b200
from pdlp_b200 import mp_model as mpm

TR = pdlp_b200.TerminationReason


def _solver_for(backend):
    import test_mp_model

    return test_mp_model._solver_for(backend)


def test_solve_with_pdlp_one_bounded_variable_no_constraints(backend):
    """ortools/linear_solver/python/model_builder_helper_test.py:104-126: maximise x on [0, 1] through an MPModelRequest
    with solver_type PDLP_LINEAR_PROGRAMMING; OPTIMAL with objective 1.  A problem with variables and no constraint."""
    request = pdlp_b200.MPModelRequest()
    variable = request.model.variable.add()
    variable.lower_bound, variable.upper_bound, variable.objective_coefficient = 0.0, 1.0, 1.0
    request.model.maximize = True
    request.solver_type = mpm.PDLP_LINEAR_PROGRAMMING
    response = pdlp_b200.pdlp_solve_proto(request, solver=_solver_for(backend))
    assert response.status == mpm.MPSOLVER_OPTIMAL
    assert response.objective_value == pytest.approx(1.0, abs=1e-7)
    assert list(response.variable_value) == pytest.approx([1.0], abs=1e-7)
    assert len(response.dual_value) == 0


def test_constraints_without_variables(backend):
    """The mirror case: rows and no column.  Feasible iff 0 lies in every row's bounds; the optimum is the offset."""
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(0, 2)
    qp.constraint_lower_bounds = [-1.0, -np.inf]
    qp.constraint_upper_bounds = [1.0, 0.0]
    qp.objective_offset = 3.0
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.iteration_limit = 1000
    out = backend.solve(qp, p)
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    assert len(out.primal_solution) == 0 and len(out.dual_solution) == 2
    for info in out.solve_log.solution_stats.convergence_information:
        if info.candidate_type == out.solve_log.solution_type:
            assert info.primal_objective == pytest.approx(3.0)


# ---- primal_dual_hybrid_gradient_test.cc:1186-1273: with a presolver in the loop, nonzeros below 1e-50 are rejected ----

def _test_lp_with_tiny(field, index, lower_index=None):
    import fixtures

    qp = fixtures.make_test_lp()
    values = np.array(getattr(qp, field), dtype=float)
    values[index] = 1.0e-60
    setattr(qp, field, values)
    if lower_index is not None:  # :1225-1235 keeps the bounds consistent: 0 <= x_1 <= 1e-60
        lower = np.array(qp.variable_lower_bounds, dtype=float)
        lower[lower_index] = 0.0
        qp.variable_lower_bounds = lower
    return qp


TINY_VALUES = {
    "constraint_lower_bound": lambda: _test_lp_with_tiny("constraint_lower_bounds", 2),
    "variable_bound": lambda: _test_lp_with_tiny("variable_upper_bounds", 1, lower_index=1),
    "objective_vector": lambda: _test_lp_with_tiny("objective_vector", 3),
}


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(TINY_VALUES))
def test_detects_excessively_small_values_with_presolve(case):
    """Detects...WithPresolve: the same problems solve to OPTIMAL without a presolver
    (test_excessively_small_values_ok_without_presolve); in the hand-off - the device path's form of use_glop - the
    original problem's statistics are checked with the small-value rule on (CheckProblemStats, pdhg.cc:731-897)."""
    qp = TINY_VALUES[case]()
    identity = lambda x, y: (x, y)
    out = pdlp_b200.primal_dual_hybrid_gradient_presolved(qp, TINY_VALUES[case](), identity, pdlp_b200.PrimalDualHybridGradientParams())
    assert out.solve_log.termination_reason == TR.TERMINATION_REASON_INVALID_PROBLEM, out.solve_log.termination_string


# ---- the boundary is re-entrant (one solve owns its own pool; here: its own stream and workspace) ----

@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_concurrent_solves_from_two_threads_match_sequential_ones():
    """Two host threads solve different LPs on the same GPU at the same time.  Every reduction in the kernels has a fixed
    order, so each result must equal, bit for bit, the one the same problem gives when solved alone."""
    import threading

    import fixtures

    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-6
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-6
    p.termination_criteria.iteration_limit = 50000
    problems = [fixtures.random_lp(300, 600, 5, 11), fixtures.random_lp(500, 900, 4, 12)]
    alone = [pdlp_b200.primal_dual_hybrid_gradient(q, p, device=0) for q in problems]
    together = [None, None]
    errors = []

    def work(k):
        try:
            for _ in range(3):  # several solves per thread, so that set-up, loop and download phases interleave
                together[k] = pdlp_b200.primal_dual_hybrid_gradient(problems[k], p, device=0)
        except Exception as e:  # noqa: BLE001 - reported by the assertion below
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for a, b in zip(alone, together):
        assert a.solve_log.termination_reason == b.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
        assert a.solve_log.iteration_count == b.solve_log.iteration_count
        np.testing.assert_array_equal(a.primal_solution, b.primal_solution)
        np.testing.assert_array_equal(a.dual_solution, b.dual_solution)


@pytest.mark.gpu
def test_plain_c_caller_solves_test_lp():
    """tests/c/caller.c (gcc -std=c99 against include/pdlp_b200.h): pdlp_b200_solve with NULL params, a callback and counters."""
    import subprocess

    import test_cabi

    run = subprocess.run([test_cabi.build_c_caller(), "gpu"], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0 and "gpu ok" in run.stdout, (run.returncode, run.stdout, run.stderr)
