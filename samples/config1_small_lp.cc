// BASELINE.json config 1 through the C++ header: the 4 x 4 LP of the reference's sample
// (ortools/pdlp/samples/simple_pdlp_program.cc; TestLp of test_util.h:33-48) under the reference's qualified names.
// Build, then run on a machine with a CUDA device:
//   g++ -std=c++17 -Iinclude samples/config1_small_lp.cc -Lor-tools_b200/lib -lpdlp_b200 -Wl,-rpath,$PWD/or-tools_b200/lib -o config1_small_lp
#include <cstdio>
#include <limits>
#include <vector>

#include "pdlp_b200/primal_dual_hybrid_gradient.h"

namespace pdlp = operations_research::pdlp;

struct Column { double cost, lower, upper; };
struct Row { double lower, upper; };

static pdlp::QuadraticProgram BuildProblem() {
  const double inf = std::numeric_limits<double>::infinity();
  const std::vector<Column> columns = {{5.5, -inf, inf}, {-2, -2, inf}, {-1, -inf, 6}, {1, 2.5, 3.5}};
  const std::vector<Row> rows = {{12, 12}, {-inf, 7}, {-4, inf}, {-1, 1}};
  const std::vector<pdlp::Triplet> entries = {{0, 0, 2}, {0, 1, 1}, {0, 2, 1}, {0, 3, 2}, {1, 0, 1}, {1, 2, 1}, {2, 0, 4}, {3, 2, 1.5}, {3, 3, -1}};
  pdlp::QuadraticProgram qp(static_cast<int64_t>(columns.size()), static_cast<int64_t>(rows.size()));
  for (size_t j = 0; j < columns.size(); ++j) {
    qp.objective_vector[j] = columns[j].cost;
    qp.variable_lower_bounds[j] = columns[j].lower;
    qp.variable_upper_bounds[j] = columns[j].upper;
  }
  for (size_t i = 0; i < rows.size(); ++i) {
    qp.constraint_lower_bounds[i] = rows[i].lower;
    qp.constraint_upper_bounds[i] = rows[i].upper;
  }
  qp.constraint_matrix.setFromTriplets(entries.begin(), entries.end());
  qp.objective_offset = -14;
  return qp;
}

int main() {
  pdlp::PrimalDualHybridGradientParams params;
  auto& criteria = params.termination_criteria.mut().simple_optimality_criteria.mut();
  criteria.eps_optimal_absolute = 1.0e-6;
  criteria.eps_optimal_relative = 1.0e-6;
  params.verbosity_level = 0;
  params.presolve_options.mut().use_glop = false;

  const pdlp::SolverResult result = pdlp::PrimalDualHybridGradient(BuildProblem(), params);
  const pdlp::SolveLog& log = result.solve_log;
  std::printf("%s after %d iterations (%.4f s); returned point: %s\n", pdlp::TerminationReason_Name(log.termination_reason),
              static_cast<int>(log.iteration_count), static_cast<double>(log.solve_time_sec), pdlp::PointType_Name(log.solution_type));
  if (const auto info = pdlp::GetConvergenceInformation(log.solution_stats.get(), log.solution_type)) {
    std::printf("objective: primal %.9g  dual %.9g\n", static_cast<double>(info->primal_objective), static_cast<double>(info->dual_objective));
    std::printf("residuals: primal %.3e  dual %.3e\n", static_cast<double>(info->l2_primal_residual), static_cast<double>(info->l2_dual_residual));
  }
  for (size_t j = 0; j < result.primal_solution.size(); ++j) std::printf("x%zu %12.6f %14.6f\n", j, result.primal_solution[j], result.reduced_costs[j]);
  for (size_t i = 0; i < result.dual_solution.size(); ++i) std::printf("y%zu %12.6f\n", i, result.dual_solution[i]);
  return log.termination_reason == pdlp::TERMINATION_REASON_OPTIMAL ? 0 : 1;
}
