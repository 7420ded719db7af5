"""BASELINE.json config 1 through the Python module: the 4 x 4 LP that the reference ships as its sample
(ortools/pdlp/samples/simple_pdlp_program.{py,cc}; the same problem as TestLp, test_util.h:33-48), solved at the
sample's tolerances and reported as one table.  Usage (needs a CUDA device; the solver has no CPU path):

    python samples/config1_small_lp.py
"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pdlp_b200  # noqa: E402

INF = math.inf
# (row, column, coefficient) of   2 x0 + x1 + x2 + 2 x3 = 12;   x0 + x2 <= 7;   4 x0 >= -4;   -1 <= 1.5 x2 - x3 <= 1
ENTRIES = [(0, 0, 2.0), (0, 1, 1.0), (0, 2, 1.0), (0, 3, 2.0), (1, 0, 1.0), (1, 2, 1.0), (2, 0, 4.0), (3, 2, 1.5), (3, 3, -1.0)]
ROW_BOUNDS = [(12.0, 12.0), (-INF, 7.0), (-4.0, INF), (-1.0, 1.0)]
COLUMNS = [  # (cost, lower, upper): minimise 5.5 x0 - 2 x1 - x2 + x3 - 14
    (5.5, -INF, INF), (-2.0, -2.0, INF), (-1.0, -INF, 6.0), (1.0, 2.5, 3.5)]
OFFSET = -14.0


def build_problem():
    qp = pdlp_b200.QuadraticProgram()
    qp.resize_and_initialize(len(COLUMNS), len(ROW_BOUNDS))
    rows, cols, values = zip(*ENTRIES)
    qp.constraint_matrix = pdlp_b200.set_matrix_from_triplets(rows, cols, values, (len(ROW_BOUNDS), len(COLUMNS)))
    qp.objective_vector = [c for c, _, _ in COLUMNS]
    qp.variable_lower_bounds = [lo for _, lo, _ in COLUMNS]
    qp.variable_upper_bounds = [hi for _, _, hi in COLUMNS]
    qp.constraint_lower_bounds = [lo for lo, _ in ROW_BOUNDS]
    qp.constraint_upper_bounds = [hi for _, hi in ROW_BOUNDS]
    qp.objective_offset = OFFSET
    return qp


def sample_params():
    params = pdlp_b200.PrimalDualHybridGradientParams()
    params.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1.0e-6
    params.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1.0e-6
    params.verbosity_level = 0
    params.presolve_options.use_glop = False
    return params


def report(result, out=sys.stdout):
    log = result.solve_log
    reason = pdlp_b200.solve_log_pb2.TerminationReason.Name(log.termination_reason)
    point = pdlp_b200.solve_log_pb2.PointType.Name(log.solution_type)
    print(f"{reason} after {log.iteration_count} iterations ({log.solve_time_sec:.4f} s); returned point: {point}", file=out)
    for info in log.solution_stats.convergence_information:
        if info.candidate_type == log.solution_type:
            print(f"objective: primal {info.primal_objective:.9g}  dual {info.dual_objective:.9g}", file=out)
            print(f"residuals: primal {info.l2_primal_residual:.3e}  dual {info.l2_dual_residual:.3e}", file=out)
    print(f"{'':>4} {'x':>12} {'reduced cost':>14}", file=out)
    for j, (x, rc) in enumerate(zip(result.primal_solution, result.reduced_costs)):
        print(f"x{j:<3} {x:12.6f} {rc:14.6f}", file=out)
    for i, y in enumerate(result.dual_solution):
        print(f"y{i:<3} {y:12.6f}", file=out)


def run(solve=pdlp_b200.primal_dual_hybrid_gradient, out=sys.stdout):
    result = solve(build_problem(), sample_params())
    report(result, out)
    return result


if __name__ == "__main__":
    ok = run().solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    sys.exit(0 if ok else 1)
