"""interrupt_solve with per-iteration semantics on the device (`-m gpu`): termination.cc:178 tests the flag at every
iteration (primal_dual_hybrid_gradient.cc:2147).  The device-resident loop reads a host-mapped copy of the flag in its
step-decision kernel, so a flag raised in the middle of a batch of blind-enqueued trial steps stops that batch."""
import threading
import time

import numpy as np
import pytest

import fixtures
import pdlp_b200

pytestmark = pytest.mark.gpu

TR = pdlp_b200.TerminationReason


def _never_converging_params():
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    # Host events only every 4096 iterations: whatever stops the solve in between is the device's own test of the flag.
    p.major_iteration_frequency = 4096
    p.termination_check_frequency = 4096
    p.termination_criteria.iteration_limit = 400000
    return p


def test_flag_raised_mid_solve_stops_the_device_loop_between_host_events():
    qp = fixtures.random_lp(2000, 4000, 5, 3)
    # CUDA context, library load and the first allocations must not eat the 0.25 s before the flag goes up.
    warm = _never_converging_params()
    warm.termination_criteria.iteration_limit = 64
    pdlp_b200.primal_dual_hybrid_gradient(qp, warm, device=0)
    def attempt(delay):
        flag = np.zeros(1, dtype=np.int32)
        raised = {}

        def raise_later():
            time.sleep(delay)
            raised["t"] = time.perf_counter()
            flag[0] = 1

        t = threading.Thread(target=raise_later)
        t.start()
        result = pdlp_b200.primal_dual_hybrid_gradient(qp, _never_converging_params(), interrupt_solve=flag, device=0)
        returned = time.perf_counter()
        t.join()
        return result.solve_log, pdlp_b200.last_solve_counters(), returned - raised["t"]

    # The flag may go up while the host happens to be at one of its own events (a millisecond every 4096 iterations, and
    # the artificial restarts at powers of two): the host then sees it first, which is correct but shows nothing about
    # the device.  Such an attempt is repeated with another delay.
    for delay in (0.25, 0.31, 0.37):
        log, counters, latency = attempt(delay)
        assert log.termination_reason == TR.TERMINATION_REASON_INTERRUPTED_BY_USER
        assert log.iteration_count < 400000
        if counters["interrupts_seen_on_device"] == 1:
            break
    assert counters["interrupts_seen_on_device"] == 1, counters
    # Not at a host event: those sit at multiples of 4096 (and at the artificial restarts of the first iterations).
    assert log.iteration_count > 64 and log.iteration_count % 4096 != 0, log.iteration_count
    # The flag is seen within one iteration; what follows is the drain of at most two 32-trial chunks, the final
    # termination check and the download of the result (a few ms at this size).  The assertions above are the proof of
    # the device-side stop; this bound only guards against a regression to batch granularity (a batch of 4096
    # iterations is well over 100 ms) without being sensitive to a loaded host.
    assert latency < 0.1, latency


def test_flag_already_raised_stops_at_iteration_zero_like_the_reference():
    flag = np.ones(1, dtype=np.int32)
    result = pdlp_b200.primal_dual_hybrid_gradient(fixtures.make_test_lp(), _never_converging_params(), interrupt_solve=flag, device=0)
    assert result.solve_log.termination_reason == TR.TERMINATION_REASON_INTERRUPTED_BY_USER
    assert result.solve_log.iteration_count == 0
    assert pdlp_b200.last_solve_counters()["interrupts_seen_on_device"] == 0  # the host's check at iteration 0 saw it first


def test_a_flag_that_stays_down_changes_nothing():
    qp = fixtures.random_lp(300, 600, 5, 1)
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-6
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-6
    flag = np.zeros(1, dtype=np.int32)
    with_flag = pdlp_b200.primal_dual_hybrid_gradient(qp, p, interrupt_solve=flag, device=0)
    without = pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=0)
    assert with_flag.solve_log.termination_reason == without.solve_log.termination_reason == TR.TERMINATION_REASON_OPTIMAL
    assert with_flag.solve_log.iteration_count == without.solve_log.iteration_count
    np.testing.assert_array_equal(with_flag.primal_solution, without.primal_solution)
