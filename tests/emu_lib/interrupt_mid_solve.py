"""TEST INFRASTRUCTURE ONLY.  Run by tests/test_full_emulation.py with PDLP_B200_LIBRARY = the emulated library: the
per-iteration interrupt (termination.cc:178) through the whole host driver - mapped flag, watcher thread, chunked
enqueue, the decision kernel's own test of the flag - with host events 4096 iterations apart, so that only the device
side can have stopped the loop where it stops.  The B200 version of this test (tests/test_gpu_interrupt.py) raises the
flag 0.25 s into the solve; an emulated iteration takes milliseconds, so this one waits 2 s.  Prints one JSON line."""
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import fixtures  # noqa: E402
import pdlp_b200  # noqa: E402


def main():
    qp = fixtures.random_lp(40, 100, 4, 21)
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 0.0
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 0.0
    p.major_iteration_frequency = 4096
    p.termination_check_frequency = 4096
    p.termination_criteria.iteration_limit = 400000
    def attempt(delay):
        flag = np.zeros(1, dtype=np.int32)
        raised = {}

        def raise_later():
            time.sleep(delay)
            raised["t"] = time.perf_counter()
            flag[0] = 1

        t = threading.Thread(target=raise_later)
        t.start()
        result = pdlp_b200.primal_dual_hybrid_gradient(qp, p, interrupt_solve=flag, device=0)
        returned = time.perf_counter()
        t.join()
        log = result.solve_log
        return {"reason": int(log.termination_reason), "iterations": int(log.iteration_count), "latency": returned - raised["t"],
                "seen_on_device": int(pdlp_b200.last_solve_counters()["interrupts_seen_on_device"]), "delay": delay}

    # The flag may go up while the host is at one of its own events (the artificial restarts of the first iterations): the
    # host then sees it first, which is correct but shows nothing about the device.  Such an attempt is repeated.
    for delay in (2.0, 2.7, 3.5):
        got = attempt(delay)
        if got["seen_on_device"] == 1:
            break
    print(json.dumps(got))


if __name__ == "__main__":
    main()
