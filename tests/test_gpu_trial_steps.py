"""`pdlp_b200_trial_steps`, the unit bench.py times, against the numpy restatement of
Solver::TakeAdaptiveStep (primal_dual_hybrid_gradient.cc:2295-2376 with :1659-1771 and
sharded_optimization_utils.cc:54-66) that tests/test_kernel_emulation.py holds the emulated kernels
to: same accept / reject sequence, iterates to 1e-9, step size to 1e-9 (`-m gpu`; first run on a B200
in round 2 as scripts/check_trial_steps_gpu.py: worst error 6.6e-13)."""
import numpy as np
import pytest

import pdlp_b200
import test_kernel_emulation as emu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,qp_case,step,w", [(2, False, 50.0, 0.2), (6, True, 50.0, 0.2), (3, False, 0.3, 1.1)])
def test_trial_steps_match_the_restated_adaptive_rule(seed, qp_case, step, w):
    K, c, q, vlb, vub, clb, cub, x0, y0 = prob = emu.step_problem(seed, qp_case)
    trials = 14
    want = emu.reference_adaptive_trials(*prob, step, w, trials)
    qp = pdlp_b200.QuadraticProgram()
    qp.constraint_matrix = K.tocsc()
    qp.objective_vector = c
    qp.variable_lower_bounds, qp.variable_upper_bounds = vlb, vub
    qp.constraint_lower_bounds, qp.constraint_upper_bounds = clb, cub
    if q is not None:
        qp.set_objective_matrix_diagonal(q)
    dp = pdlp_b200.DeviceProblem(qp)
    try:
        x, y, st, _, _ = dp.trial_steps(trials, step, w, x0, y0)
    finally:
        dp.close()
    assert (st["iterations_completed"], st["num_rejected_steps"]) == (want["completed"], want["rejected"])
    assert st["iterations_completed"] + st["num_rejected_steps"] + st["pending_rejected_steps"] == trials == st["total_trials"]
    assert not st["forced_numerical_termination"]
    np.testing.assert_allclose(x, want["x"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(y, want["y"], rtol=1e-9, atol=1e-9)
    assert abs(st["step_size"] - want["step"]) <= 1e-9 * want["step"]
