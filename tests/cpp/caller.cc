// A caller of the reference's public C++ API, written against include/pdlp_b200/primal_dual_hybrid_gradient.h
// under the reference's own qualified names (ortools/pdlp/primal_dual_hybrid_gradient.h:140-156).
// Built and run by tests/test_cpp_header.py.   usage: caller cpu | caller gpu
#include <cmath>
#include <cstdio>
#include <cstring>

#include "pdlp_b200/primal_dual_hybrid_gradient.h"

namespace pdlp = operations_research::pdlp;

// TestLp of ortools/pdlp/test_util.cc:35-60: optimum -34 at x = [-1, 8, 1, 2.5].
static pdlp::QuadraticProgram TestLp() {
  const double inf = std::numeric_limits<double>::infinity();
  pdlp::QuadraticProgram qp;
  qp.ResizeAndInitialize(4, 4);
  qp.constraint_matrix.outer_index = {0, 3, 4, 7, 9};
  qp.constraint_matrix.inner_index = {0, 1, 2, 0, 0, 1, 3, 0, 3};
  qp.constraint_matrix.values = {2, 1, 4, 1, 1, 1, 1.5, 2, -1};
  qp.objective_vector = {5.5, -2, -1, 1};
  qp.constraint_lower_bounds = {12, -inf, -4, -1};
  qp.constraint_upper_bounds = {12, 7, inf, 1};
  qp.variable_lower_bounds = {-inf, -2, -inf, 2.5};
  qp.variable_upper_bounds = {inf, inf, 6, 3.5};
  qp.objective_offset = -14;
  return qp;
}

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  pdlp::PrimalDualHybridGradientParams params;
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_absolute = 1.0e-6;
  params.termination_criteria.mut().simple_optimality_criteria.mut().eps_optimal_relative = 1.0e-6;
  if (!gpu) {
    // Errors found before the device is touched come back the reference's way: a normal return with the
    // reason in the log (primal_dual_hybrid_gradient.cc:717-725, :2787-2831).
    pdlp::QuadraticProgram bad = TestLp();
    bad.constraint_lower_bounds.pop_back();
    const pdlp::SolverResult r1 = pdlp::PrimalDualHybridGradient(bad, params);
    if (r1.solve_log.termination_reason != pdlp::TERMINATION_REASON_INVALID_PROBLEM) return 1;
    pdlp::PrimalDualHybridGradientParams bad_params = params;
    bad_params.major_iteration_frequency = 0;
    const pdlp::SolverResult r2 = pdlp::PrimalDualHybridGradient(TestLp(), bad_params);
    if (r2.solve_log.termination_reason != pdlp::TERMINATION_REASON_INVALID_PARAMETER) return 2;
    pdlp::PrimalAndDualSolution start{{0, 0, 0}, {0, 0, 0, 0}};  // wrong primal size
    const pdlp::SolverResult r3 = pdlp::PrimalDualHybridGradient(TestLp(), params, start);
    if (r3.solve_log.termination_reason != pdlp::TERMINATION_REASON_INVALID_INITIAL_SOLUTION) return 3;
    if (!pdlp::IsLinearProgram(TestLp()) || !pdlp::HasValidBounds(TestLp()) || !pdlp::ValidateQuadraticProgramDimensions(TestLp()).empty()) return 4;
    // HasValidBoundsTest (quadratic_program_test.cc:134-148): a constraint and a variable with lower > upper bound.
    pdlp::QuadraticProgram crossed_row = TestLp(), crossed_column = TestLp();
    crossed_row.constraint_lower_bounds[1] = 8;  // upper bound 7
    crossed_column.variable_lower_bounds[3] = 4;  // upper bound 3.5
    if (pdlp::HasValidBounds(crossed_row) || pdlp::HasValidBounds(crossed_column)) return 7;
    // internal::ComputeStatuses (primal_dual_hybrid_gradient.h:164; ComputeStatusesTest.AtOptimum, ..._test.cc:2076-2092).
    const pdlp::ProblemStatuses st = pdlp::internal::ComputeStatuses(TestLp(), pdlp::PrimalAndDualSolution{{-1, 8, 1, 2.5}, {-2, 0, 2.375, 2.0 / 3}});
    using CS = pdlp::ConstraintStatus;
    using VS = pdlp::VariableStatus;
    if (st.constraint_statuses != std::vector<CS>{CS::FIXED_VALUE, CS::BASIC, CS::AT_LOWER_BOUND, CS::AT_LOWER_BOUND}) return 5;
    if (st.variable_statuses != std::vector<VS>{VS::BASIC, VS::BASIC, VS::BASIC, VS::AT_LOWER_BOUND}) return 6;
    // CombineRepeatedTripletsInPlace / SetEigenMatrixFromTriplets (quadratic_program_test.cc:426-487).
    std::vector<pdlp::Triplet> t1 = {{1, 2, 3.0}, {1, 2, -1.0}, {1, 1, 0.0}};
    pdlp::internal::CombineRepeatedTripletsInPlace(t1);
    if (t1.size() != 2 || t1[0].row() != 1 || t1[0].col() != 2 || t1[0].value() != 2.0 || t1[1].col() != 1 || t1[1].value() != 0.0) return 8;
    std::vector<pdlp::Triplet> t2 = {{1, 2, 3.0}, {1, 2, 1.0}, {1, 2, 2.0}}, t3;
    pdlp::internal::CombineRepeatedTripletsInPlace(t2);
    pdlp::internal::CombineRepeatedTripletsInPlace(t3);
    if (t2.size() != 1 || t2[0].value() != 6.0 || !t3.empty()) return 9;
    pdlp::SparseMatrixCsc tiny;
    tiny.rows = tiny.cols = 2;
    pdlp::SetEigenMatrixFromTriplets({{0, 0, 1.0}, {1, 0, -1.0}, {0, 0, 0.0}, {1, 1, 1.0}, {0, 0, 1.0}}, tiny);  // dense [[2, 0], [-1, 1]]
    if (tiny.outer_index != std::vector<int64_t>{0, 2, 3} || tiny.inner_index != std::vector<int64_t>{0, 1, 1} || tiny.values != std::vector<double>{2, -1, 1}) return 13;
    pdlp::SparseMatrixCsc none;
    none.rows = none.cols = 2;
    pdlp::SetEigenMatrixFromTriplets({}, none);
    if (none.outer_index != std::vector<int64_t>{0, 0, 0} || none.nonZeros() != 0) return 14;
    // The sample's way of building TestLp gives the same arrays as the ones typed in above.
    pdlp::QuadraticProgram built(4, 4);
    const std::vector<pdlp::Triplet> entries = {{0, 0, 2}, {0, 1, 1}, {0, 2, 1}, {0, 3, 2}, {1, 0, 1}, {1, 2, 1}, {2, 0, 4}, {3, 2, 1.5}, {3, 3, -1}};
    built.constraint_matrix.setFromTriplets(entries.begin(), entries.end());
    const pdlp::QuadraticProgram typed = TestLp();
    if (built.constraint_matrix.outer_index != typed.constraint_matrix.outer_index || built.constraint_matrix.inner_index != typed.constraint_matrix.inner_index ||
        built.constraint_matrix.values != typed.constraint_matrix.values || built.variable_lower_bounds.size() != 4) return 15;
    // GetConvergenceInformation / GetInfeasibilityInformation / GetPointMetadata, the "GetsCorrectEntry" cases of
    // iteration_stats_test.cc:256-336: two entries (current = 1, average = 2), none for the iterate difference.
    pdlp::IterationStats stats;
    const int kinds[2] = {pdlp::POINT_TYPE_CURRENT_ITERATE, pdlp::POINT_TYPE_AVERAGE_ITERATE};
    for (int k = 0; k < 2; ++k) {
      stats.convergence_information.emplace_back();
      stats.convergence_information.back().candidate_type = kinds[k];
      stats.convergence_information.back().primal_objective = k + 1.0;
      stats.infeasibility_information.emplace_back();
      stats.infeasibility_information.back().candidate_type = kinds[k];
      stats.infeasibility_information.back().primal_ray_linear_objective = k + 1.0;
      stats.point_metadata.emplace_back();
      stats.point_metadata.back().point_type = kinds[k];
      stats.point_metadata.back().active_primal_variable_count = k + 1;
    }
    for (int k = 0; k < 2; ++k) {
      const auto ci = pdlp::GetConvergenceInformation(stats, kinds[k]);
      if (!ci.has_value() || ci->candidate_type != kinds[k] || ci->primal_objective != k + 1.0) return 16;
      const auto ii = pdlp::GetInfeasibilityInformation(stats, kinds[k]);
      if (!ii.has_value() || ii->candidate_type != kinds[k] || ii->primal_ray_linear_objective != k + 1.0) return 17;
      const auto pm = pdlp::GetPointMetadata(stats, kinds[k]);
      if (!pm.has_value() || pm->point_type != kinds[k] || pm->active_primal_variable_count != k + 1) return 18;
    }
    if (pdlp::GetConvergenceInformation(stats, pdlp::POINT_TYPE_ITERATE_DIFFERENCE).has_value() ||
        pdlp::GetInfeasibilityInformation(stats, pdlp::POINT_TYPE_ITERATE_DIFFERENCE).has_value() ||
        pdlp::GetPointMetadata(stats, pdlp::POINT_TYPE_ITERATE_DIFFERENCE).has_value()) return 20;
    if (std::strcmp(pdlp::TerminationReason_Name(pdlp::TERMINATION_REASON_OPTIMAL), "TERMINATION_REASON_OPTIMAL") != 0) return 19;
    std::printf("cpu ok: %s | %s\n", r1.solve_log.termination_string.get().c_str(), r2.solve_log.termination_string.get().c_str());
    return 0;
  }
  int callbacks = 0;
  std::atomic<bool> interrupt{false};
  const pdlp::SolverResult r = pdlp::PrimalDualHybridGradient(TestLp(), params, &interrupt, [&](const pdlp::IterationCallbackInfo&) { ++callbacks; });
  if (r.solve_log.termination_reason != pdlp::TERMINATION_REASON_OPTIMAL) return 10;
  const double want[4] = {-1, 8, 1, 2.5};
  for (int j = 0; j < 4; ++j) if (std::fabs(r.primal_solution[j] - want[j]) > 1.0e-4) return 11;
  double objective = -14;
  const double c[4] = {5.5, -2, -1, 1};
  for (int j = 0; j < 4; ++j) objective += c[j] * r.primal_solution[j];
  if (std::fabs(objective + 34.0) > 1.0e-4 || callbacks < 2) return 12;
  std::printf("gpu ok: OPTIMAL in %d iterations, objective %.6f, %d callbacks, %lld kernel launches\n", static_cast<int>(r.solve_log.iteration_count), objective,
              callbacks, pdlp::LastSolveCounters().kernel_launches);
  return 0;
}
