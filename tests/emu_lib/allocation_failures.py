"""TEST INFRASTRUCTURE ONLY.  Run by tests/test_full_emulation.py with PDLP_B200_LIBRARY = the emulated library: every
device allocation of a solve fails once, in turn (the emulated cudaMalloc's fault injection).  Each attempt must come back
as the ABI's CUDA-error status with a message - no crash, no hang - and the solve after it must be unaffected."""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import fixtures  # noqa: E402
import pdlp_b200  # noqa: E402


def main():
    countdown = ctypes.c_longlong.in_dll(pdlp_b200._lib, "_ZZ24pdlp_emu_fail_allocationE9countdown")
    qp = fixtures.make_test_lp()
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.iteration_limit = 40
    reference = pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=0)
    failed, messages, k = 0, set(), 0
    while True:
        k += 1
        countdown.value = k
        try:
            got = pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=0)
        except RuntimeError as e:
            failed += 1
            messages.add(str(e)[:60])
            assert countdown.value == 0
            continue
        # the k-th allocation does not exist any more: the solve ran to its end untouched
        assert countdown.value > 0
        countdown.value = 0
        assert np.array_equal(got.primal_solution, reference.primal_solution) and got.solve_log.iteration_count == reference.solve_log.iteration_count
        break
    after = pdlp_b200.primal_dual_hybrid_gradient(qp, p, device=0)
    print(json.dumps({"allocations_per_solve": k - 1, "failed_cleanly": failed, "messages": sorted(messages),
                      "solve_after_failures_identical": bool(np.array_equal(after.primal_solution, reference.primal_solution))}))


if __name__ == "__main__":
    main()
