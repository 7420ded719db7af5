#!/usr/bin/env python
"""GPU check (not yet a pytest case: written after the round's GPU budget was spent, promote it to
tests/ once it has run green on a B200): the bench's timed unit, pdlp_b200_trial_steps, against the
numpy restatement of Solver::TakeAdaptiveStep that tests/test_kernel_emulation.py already holds the
emulated kernels to.

    gpurun -- python scripts/check_trial_steps_gpu.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest  # noqa: E402,F401  (puts the package on sys.path the way the tests do)
import pdlp_b200  # noqa: E402
import test_kernel_emulation as emu  # noqa: E402


def main():
    worst = 0.0
    for seed, qp_case, step, w in ((2, False, 50.0, 0.2), (6, True, 50.0, 0.2), (3, False, 0.3, 1.1)):
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
        x, y, st, _, _ = dp.trial_steps(trials, step, w, x0, y0)
        dp.close()
        ok_counts = (st["iterations_completed"], st["num_rejected_steps"]) == (want["completed"], want["rejected"])
        ex = np.max(np.abs(x - want["x"]) / (1e-12 + np.abs(want["x"])).clip(1.0))
        ey = np.max(np.abs(y - want["y"]) / (1e-12 + np.abs(want["y"])).clip(1.0))
        es = abs(st["step_size"] - want["step"]) / want["step"]
        worst = max(worst, ex, ey, es)
        print(f"seed {seed} qp {qp_case}: counts {'ok' if ok_counts else 'DIFFER'} ({st['iterations_completed']}/{st['num_rejected_steps']} vs "
              f"{want['completed']}/{want['rejected']}), max err x {ex:.2e} y {ey:.2e} step {es:.2e}")
        if not ok_counts:
            return 1
    print("worst error", worst)
    return 0 if worst < 1e-9 else 1


if __name__ == "__main__":
    sys.exit(main())
