"""One launch of every kernel family of the PDHG path on a resident problem, for an
`ncu --set full` capture (B200_PROFILING.md): set-up (transpose, stats, Ruiz / L2 rescaling),
three trial steps (primal step, K x + dual step, K^T y + step test, decision, averaging), one
termination check (KKT row / column passes), one restart evaluation (Lagrangian gradients + the
trust-region root find), the reduced costs, and a 70-iteration solve (host events with their products pass).

    python scripts/profile_kernels.py [workload=c3] [scale=1.0]
    ncu --set full --clock-control none --import-source on -k regex:pdlp_b200 -c 120 -o gpurun_out/prof_all python scripts/profile_kernels.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import bench  # noqa: E402
import pdlp_b200  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    qp = bench.make_problem(name, scale)
    m, n = qp.constraint_matrix.shape
    dp = pdlp_b200.DeviceProblem(qp)          # upload + CSR build
    stats = dp.stats()                        # InfoOp / RowNormOp reductions
    dp.rescale(1, True)                       # one Ruiz pass + the L2 pass + matrix / vector scaling
    stats = dp.stats()
    step = 1.0 / max(1e-20, stats.constraint_matrix_abs_max)
    weight = stats.objective_vector_l2_norm / stats.combined_bounds_l2_norm
    x, y, st, seconds, launches = dp.trial_steps(3, step, weight)
    p = bench.solve_params(name)
    k = dp.kkt_stats(p, x, y)                 # KktRowOp / KktColOp
    b = dp.localized_bounds(x, y, weight, 1.0)  # DualGradOp, PrimalGradMapOp, trust-region passes
    rc = dp.reduced_costs(p, x, y)
    dp.close()
    # A short solve: the host events' products pass (rowstream_multi_kernel), the dense KKT / gradient passes on it
    # and the trust-region passes, as the solver launches them.
    p.termination_criteria.iteration_limit = 70
    r = pdlp_b200.primal_dual_hybrid_gradient(qp, p)
    print(f"short solve: {r.solve_log.iteration_count} iterations, reason {r.solve_log.termination_reason}")
    print(f"{name}: n={n} m={m} nnz={qp.constraint_matrix.nnz}  3 steps in {seconds * 1e3:.3f} ms, {launches} launches; "
          f"gap {k.convergence_information[0].primal_objective - k.convergence_information[0].dual_objective:.4g}, bounds {b}, |rc| {np.abs(rc).max():.4g}")


if __name__ == "__main__":
    main()
