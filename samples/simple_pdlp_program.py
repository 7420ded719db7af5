"""Solves a small LP through pdlp_b200's direct Python API: the program of
ortools/pdlp/samples/simple_pdlp_program.py with the module name changed and nothing else (BASELINE.json's
config 1: the 4 x 4 LP of samples/simple_pdlp_program.cc:49-65, objective -34 at x = [-1, 8, 1, 2.5]).

    python samples/simple_pdlp_program.py          # needs a CUDA device: the solver has no CPU path
"""
import os
import sys

import numpy as np
import scipy.sparse

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pdlp_b200 as pdlp  # noqa: E402  (stands in for ortools.pdlp.python.pdlp)
from pdlp_b200 import solve_log_pb2, solvers_pb2  # noqa: E402


def simple_lp():
    """min 5.5 x_0 - 2 x_1 - x_2 + x_3 - 14 s.t.
        2 x_0 +     x_1 +   x_2 + 2 x_3  = 12
          x_0 +             x_2          <=  7
        4 x_0                            >=  -4
       -1 <=            1.5 x_2 -   x_3  <= 1
      x_0 free, -2 <= x_1, x_2 <= 6, 2.5 <= x_3 <= 3.5
    """
    lp = pdlp.QuadraticProgram()
    lp.objective_offset = -14
    lp.objective_vector = [5.5, -2, -1, 1]
    lp.constraint_lower_bounds = [12, -np.inf, -4, -1]
    lp.constraint_upper_bounds = [12, 7, np.inf, 1]
    lp.variable_lower_bounds = [-np.inf, -2, -np.inf, 2.5]
    lp.variable_upper_bounds = [np.inf, np.inf, 6, 3.5]
    lp.constraint_matrix = scipy.sparse.csc_matrix(np.array([[2, 1, 1, 2], [1, 0, 1, 0], [4, 0, 0, 0], [0, 0, 1.5, -1]]))
    return lp


def main(solve=pdlp.primal_dual_hybrid_gradient):
    params = solvers_pb2.PrimalDualHybridGradientParams()
    optimality_criteria = params.termination_criteria.simple_optimality_criteria
    optimality_criteria.eps_optimal_relative = 1.0e-6
    optimality_criteria.eps_optimal_absolute = 1.0e-6
    params.termination_criteria.time_sec_limit = np.inf
    params.num_threads = 1  # accepted and ignored: the device path has no thread pool
    params.verbosity_level = 0
    params.presolve_options.use_glop = False

    result = solve(simple_lp(), params)
    solve_log = result.solve_log

    if solve_log.termination_reason == solve_log_pb2.TERMINATION_REASON_OPTIMAL:
        print("Solve successful")
    else:
        print("Solve not successful. Status:", solve_log_pb2.TerminationReason.Name(solve_log.termination_reason))

    # The vectors are always returned; what they mean depends on termination_reason
    # (include/pdlp_b200/primal_dual_hybrid_gradient.h, as primal_dual_hybrid_gradient.h:35-58).
    print("Primal solution:", result.primal_solution)
    print("Dual solution:", result.dual_solution)
    print("Reduced costs:", result.reduced_costs)

    solution_type = solve_log.solution_type
    print("Solution type:", solve_log_pb2.PointType.Name(solution_type))
    for ci in solve_log.solution_stats.convergence_information:
        if ci.candidate_type == solution_type:
            print("Primal objective:", ci.primal_objective)
            print("Dual objective:", ci.dual_objective)

    print("Iterations:", solve_log.iteration_count)
    print("Solve time (sec):", solve_log.solve_time_sec)
    return result


if __name__ == "__main__":
    main()
