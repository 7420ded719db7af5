// Solves a small LP through the C++ entry point: the program of ortools/pdlp/samples/simple_pdlp_program.cc
// written against include/pdlp_b200/primal_dual_hybrid_gradient.h.  What changes for a caller of the reference:
// the header path, std::vector assignment instead of Eigen's "<<", and fields instead of protobuf accessors
// (x.mut() where the reference says mutable_x()).
//
// Build (then run on a machine with a CUDA device):
//   g++ -std=c++17 -Iinclude samples/simple_pdlp_program.cc -Lor-tools_b200/lib -lpdlp_b200 -Wl,-rpath,$PWD/or-tools_b200/lib -o simple_pdlp_program
#include <cstdint>
#include <iostream>
#include <limits>
#include <optional>
#include <vector>

#include "pdlp_b200/primal_dual_hybrid_gradient.h"

namespace pdlp = ::operations_research::pdlp;

constexpr double kInfinity = std::numeric_limits<double>::infinity();

// min 5.5 x_0 - 2 x_1 - x_2 +   x_3 - 14 s.t.
//     2 x_0 +     x_1 +   x_2 + 2 x_3  = 12
//       x_0 +             x_2          <=  7
//     4 x_0                            >=  -4
//    -1 <=            1.5 x_2 -   x_3  <= 1
//   x_0 free, -2 <= x_1, x_2 <= 6, 2.5 <= x_3 <= 3.5
pdlp::QuadraticProgram SimpleLp() {
  pdlp::QuadraticProgram lp(4, 4);
  lp.constraint_lower_bounds = {12, -kInfinity, -4, -1};
  lp.constraint_upper_bounds = {12, 7, kInfinity, 1};
  lp.variable_lower_bounds = {-kInfinity, -2, -kInfinity, 2.5};
  lp.variable_upper_bounds = {kInfinity, kInfinity, 6, 3.5};
  const std::vector<pdlp::Triplet> constraint_matrix_triplets = {{0, 0, 2}, {0, 1, 1}, {0, 2, 1}, {0, 3, 2}, {1, 0, 1},
                                                                 {1, 2, 1}, {2, 0, 4}, {3, 2, 1.5}, {3, 3, -1}};
  lp.constraint_matrix.setFromTriplets(constraint_matrix_triplets.begin(), constraint_matrix_triplets.end());
  lp.objective_vector = {5.5, -2, -1, 1};
  lp.objective_offset = -14;
  return lp;
}

static void Print(const char* title, const std::vector<double>& v) {
  std::cout << title << "\n";
  for (double e : v) std::cout << e << "\n";
}

int main() {
  pdlp::PrimalDualHybridGradientParams params;
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_relative = 1.0e-6;
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_absolute = 1.0e-6;
  params.termination_criteria.mut().time_sec_limit = kInfinity;
  params.num_threads = 1;  // accepted and ignored: the device path has no thread pool
  params.verbosity_level = 0;
  params.presolve_options.mut().use_glop = false;

  const pdlp::SolverResult result = pdlp::PrimalDualHybridGradient(SimpleLp(), params);
  const pdlp::SolveLog& solve_log = result.solve_log;

  if (solve_log.termination_reason == pdlp::TERMINATION_REASON_OPTIMAL) {
    std::cout << "Solve successful" << std::endl;
  } else {
    std::cout << "Solve not successful. Status: " << pdlp::TerminationReason_Name(solve_log.termination_reason) << std::endl;
  }

  // The vectors are always returned; what they mean depends on termination_reason (primal_dual_hybrid_gradient.h:35-58).
  Print("Primal solution:", result.primal_solution);
  Print("Dual solution:", result.dual_solution);
  Print("Reduced costs:", result.reduced_costs);

  const int solution_type = solve_log.solution_type;
  std::cout << "Solution type: " << pdlp::PointType_Name(solution_type) << std::endl;
  const std::optional<pdlp::ConvergenceInformation> ci = pdlp::GetConvergenceInformation(solve_log.solution_stats.get(), solution_type);
  if (ci.has_value()) {
    std::cout << "Primal objective: " << ci->primal_objective << std::endl;
    std::cout << "Dual objective: " << ci->dual_objective << std::endl;
  }

  std::cout << "Iterations: " << solve_log.iteration_count << std::endl;
  std::cout << "Solve time (sec): " << solve_log.solve_time_sec << std::endl;
  return solve_log.termination_reason == pdlp::TERMINATION_REASON_OPTIMAL ? 0 : 1;
}
