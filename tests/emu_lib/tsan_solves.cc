// TEST INFRASTRUCTURE ONLY.  Whole solves on the emulated library built in THREAD mode under ThreadSanitizer
// (tests/emu_lib/build.py --tsan: one host thread per CUDA thread, barriers as std::barrier): the stand-in for
// compute-sanitizer's racecheck.  Every kernel a solve launches - matrix build, statistics, rescaling, the three step
// rules with their decision kernels, KKT passes, restart bounds, feasibility polishing, the QP variants - runs with
// every pair of accesses of a CTA's threads visible to the sanitizer.  scripts/emulated_tsan.sh builds and runs it.
#include <cmath>
#include <cstdio>
#include <limits>
#include <vector>

#include "pdlp_b200/primal_dual_hybrid_gradient.h"

namespace pdlp = operations_research::pdlp;
static const double kInf = std::numeric_limits<double>::infinity();

static pdlp::QuadraticProgram Make(int n, int m, const std::vector<pdlp::Triplet>& entries, std::vector<double> c, std::vector<double> clb, std::vector<double> cub,
                                   std::vector<double> vlb, std::vector<double> vub, double offset) {
  pdlp::QuadraticProgram qp(n, m);
  qp.constraint_matrix.setFromTriplets(entries.begin(), entries.end());
  qp.objective_vector = c; qp.constraint_lower_bounds = clb; qp.constraint_upper_bounds = cub;
  qp.variable_lower_bounds = vlb; qp.variable_upper_bounds = vub; qp.objective_offset = offset;
  return qp;
}
static pdlp::QuadraticProgram TestLp() {  // test_util.cc:35-60, optimum -34
  return Make(4, 4, {{0, 0, 2}, {0, 1, 1}, {0, 2, 1}, {0, 3, 2}, {1, 0, 1}, {1, 2, 1}, {2, 0, 4}, {3, 2, 1.5}, {3, 3, -1}}, {5.5, -2, -1, 1}, {12, -kInf, -4, -1},
              {12, 7, kInf, 1}, {-kInf, -2, -kInf, 2.5}, {kInf, kInf, 6, 3.5}, -14);
}
static pdlp::QuadraticProgram TinyLp() {  // optimum -1
  return Make(4, 3, {{0, 0, 2}, {0, 1, 1}, {0, 2, 1}, {0, 3, 2}, {1, 0, 1}, {1, 2, 1}, {2, 2, 1}, {2, 3, -1}}, {5, 2, 1, 1}, {12, 7, 1}, {12, kInf, kInf}, {0, 0, 0, 0},
              {2, 4, 6, 3}, -14);
}
static pdlp::QuadraticProgram DiagonalQp1() {  // optimum 6
  pdlp::QuadraticProgram qp = Make(2, 1, {{0, 0, 1}, {0, 1, 1}}, {-1, -1}, {-kInf}, {1}, {1, -2}, {2, 4}, 5);
  qp.objective_matrix = std::vector<double>{4.0, 1.0};
  return qp;
}

static int Solve(const char* name, pdlp::QuadraticProgram qp, pdlp::PrimalDualHybridGradientParams params, double want) {
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_absolute = 1.0e-6;
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_relative = 1.0e-6;
  params.termination_criteria.mut().iteration_limit = 2000;
  const pdlp::SolverResult r = pdlp::PrimalDualHybridGradient(qp, params);
  const auto info = pdlp::GetConvergenceInformation(r.solve_log.solution_stats.get(), r.solve_log.solution_type);
  const double objective = info.has_value() ? static_cast<double>(info->primal_objective) : NAN;
  const bool ok = r.solve_log.termination_reason == pdlp::TERMINATION_REASON_OPTIMAL && std::fabs(objective - want) < 1e-3;
  std::printf("%-28s %s after %d iterations, objective %.6f, %lld kernel launches\n", name, pdlp::TerminationReason_Name(r.solve_log.termination_reason),
              static_cast<int>(r.solve_log.iteration_count), objective, pdlp::LastSolveCounters().kernel_launches);
  std::fflush(stdout);
  return ok ? 0 : 1;
}

int main() {
  int bad = 0;
  pdlp::PrimalDualHybridGradientParams adaptive, mp, constant, polishing, qp, unscaled;
  mp.linesearch_rule = pdlp::MALITSKY_POCK_LINESEARCH_RULE;
  constant.linesearch_rule = pdlp::CONSTANT_STEP_SIZE_RULE;
  polishing.use_feasibility_polishing = true;
  polishing.handle_some_primal_gradients_on_finite_bounds_as_residuals = false;
  unscaled.l_inf_ruiz_iterations = 0;
  unscaled.l2_norm_rescaling = false;
  unscaled.restart_strategy = pdlp::ADAPTIVE_DISTANCE_BASED;
  bad += Solve("TestLp adaptive", TestLp(), adaptive, -34.0);
  bad += Solve("TestLp Malitsky-Pock", TestLp(), mp, -34.0);
  bad += Solve("TestLp constant step", TestLp(), constant, -34.0);
  bad += Solve("TestLp unscaled, distance", TestLp(), unscaled, -34.0);
  bad += Solve("TinyLp with polishing", TinyLp(), polishing, -1.0);
  bad += Solve("DiagonalQp1", DiagonalQp1(), qp, 6.0);
  std::printf(bad == 0 ? "all solves optimal\n" : "%d solve(s) not optimal\n", bad);
  return bad;
}
