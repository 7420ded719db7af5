// The C++ entry point of the B200 PDLP solver (public header: include/pdlp_b200/, next to the two
// headers it needs, pdlp_schema.h and pdlp_termination.h; link against libpdlp_b200.so).  The
// declarations live in namespace pdlp_b200 and are also reachable under the reference's own
// qualified names, operations_research::pdlp::PrimalDualHybridGradient etc. (end of this file), so a
// caller of ortools/pdlp/primal_dual_hybrid_gradient.h compiles against this header unchanged apart
// from the container types.  tests/test_cpp_header.py compiles and runs such a caller.
// Same names, argument order, ownership
// and error convention as the reference's public header
//   ortools/pdlp/primal_dual_hybrid_gradient.h:30-156 and quadratic_program.h:53-143,
// with std::vector / a plain CSC struct where the reference uses Eigen (Eigen is not in
// this image; the CSC arrays are layout-compatible with a compressed
// Eigen::SparseMatrix<double, ColMajor, int64_t>: valuePtr / innerIndexPtr / outerIndexPtr).
//
// Every failure is reported the reference's way: a normal return whose
// solve_log.termination_reason is INVALID_PROBLEM / INVALID_PARAMETER /
// INVALID_INITIAL_SOLUTION with termination_string set.  Only CUDA-level failures
// (no device, out of memory) throw std::runtime_error: there is no CPU fallback.
#ifndef PDLP_B200_PRIMAL_DUAL_HYBRID_GRADIENT_H_
#define PDLP_B200_PRIMAL_DUAL_HYBRID_GRADIENT_H_

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <functional>
#include <limits>
#include <optional>
#include <string>
#include <vector>

#include "pdlp_schema.h"
#include "pdlp_termination.h"

namespace pdlp_b200 {

using VectorXd = std::vector<double>;

// One (row, column, value) entry: the role of Eigen::Triplet<double, int64_t> in quadratic_program.h:194-213.
struct Triplet {
  Triplet() = default;
  Triplet(int64_t row, int64_t col, double value) : row_(row), col_(col), value_(value) {}
  int64_t row() const { return row_; }
  int64_t col() const { return col_; }
  double value() const { return value_; }
 private:
  int64_t row_ = 0, col_ = 0;
  double value_ = 0.0;
};

namespace internal {
// quadratic_program.cc:346-364: sums consecutive entries with equal row and column, in place.
inline void CombineRepeatedTripletsInPlace(std::vector<Triplet>& triplets) {
  if (triplets.empty()) return;
  size_t out = 0;
  for (size_t p = 1; p < triplets.size(); ++p) {
    if (triplets[out].row() == triplets[p].row() && triplets[out].col() == triplets[p].col()) {
      triplets[out] = Triplet(triplets[out].row(), triplets[out].col(), triplets[out].value() + triplets[p].value());
    } else {
      triplets[++out] = triplets[p];
    }
  }
  triplets.resize(out + 1);
}
}  // namespace internal

// Compressed sparse column matrix, rows x cols.
struct SparseMatrixCsc {
  int64_t rows = 0, cols = 0;
  std::vector<int64_t> outer_index{0};  // cols + 1
  std::vector<int64_t> inner_index;     // row of every nonzero
  std::vector<double> values;
  int64_t nonZeros() const { return static_cast<int64_t>(values.size()); }
  // What Eigen's setFromTriplets(begin, end) leaves behind: compressed columns, rows ascending, repeated entries summed.
  template <typename It> void setFromTriplets(It begin, It end);
};
// SetEigenMatrixFromTriplets (quadratic_program.cc:315-343): `matrix` keeps its rows x cols; entries sorted by
// (column, row), repeated ones summed, an entry that sums to zero stays stored.
inline void SetEigenMatrixFromTriplets(std::vector<Triplet> triplets, SparseMatrixCsc& matrix) {
  std::stable_sort(triplets.begin(), triplets.end(), [](const Triplet& a, const Triplet& b) {
    return a.col() != b.col() ? a.col() < b.col() : a.row() < b.row();
  });
  internal::CombineRepeatedTripletsInPlace(triplets);
  matrix.outer_index.assign(matrix.cols + 1, 0);
  matrix.inner_index.clear();
  matrix.values.clear();
  for (const Triplet& t : triplets) {
    ++matrix.outer_index[t.col() + 1];
    matrix.inner_index.push_back(t.row());
    matrix.values.push_back(t.value());
  }
  for (int64_t j = 0; j < matrix.cols; ++j) matrix.outer_index[j + 1] += matrix.outer_index[j];
}
template <typename It> void SparseMatrixCsc::setFromTriplets(It begin, It end) {
  SetEigenMatrixFromTriplets(std::vector<Triplet>(begin, end), *this);
}

// quadratic_program.h:53-143.
struct QuadraticProgram {
  QuadraticProgram() = default;
  QuadraticProgram(int64_t num_variables, int64_t num_constraints) { ResizeAndInitialize(num_variables, num_constraints); }  // quadratic_program.h:55-57
  VectorXd objective_vector;
  std::optional<VectorXd> objective_matrix;  // diagonal; the 1/2 is implicit
  SparseMatrixCsc constraint_matrix;
  VectorXd constraint_lower_bounds, constraint_upper_bounds;
  VectorXd variable_lower_bounds, variable_upper_bounds;
  std::optional<std::string> problem_name;
  std::optional<std::vector<std::string>> variable_names, constraint_names;
  double objective_offset = 0.0;
  double objective_scaling_factor = 1.0;
  double ApplyObjectiveScalingAndOffset(double objective) const { return objective_scaling_factor * (objective + objective_offset); }
  void ResizeAndInitialize(int64_t num_variables, int64_t num_constraints) {
    const double inf = std::numeric_limits<double>::infinity();
    objective_vector.assign(num_variables, 0.0);
    variable_lower_bounds.assign(num_variables, -inf);
    variable_upper_bounds.assign(num_variables, inf);
    constraint_lower_bounds.assign(num_constraints, -inf);
    constraint_upper_bounds.assign(num_constraints, inf);
    constraint_matrix = SparseMatrixCsc();
    constraint_matrix.rows = num_constraints;
    constraint_matrix.cols = num_variables;
    constraint_matrix.outer_index.assign(num_variables + 1, 0);
  }
};
inline bool IsLinearProgram(const QuadraticProgram& qp) { return !qp.objective_matrix.has_value(); }
// quadratic_program.cc:37-80; "" when consistent.
std::string ValidateQuadraticProgramDimensions(const QuadraticProgram& qp);
// quadratic_program.cc:82-90.
bool HasValidBounds(const QuadraticProgram& qp);

// primal_dual_hybrid_gradient.h:30-93.
struct PrimalAndDualSolution {
  VectorXd primal_solution;
  VectorXd dual_solution;
};
struct SolverResult {
  VectorXd primal_solution;
  VectorXd dual_solution;
  VectorXd reduced_costs;
  SolveLog solve_log;
};
enum class IterationType {
  kNormal = 0,
  kPrimalFeasibility = 1,
  kDualFeasibility = 2,
  kPresolveTermination = 3,  // only from the presolve hand-off (pdlp_b200_solve_presolved) when the caller's presolver solved the problem
  kNormalTermination = 4,
  kFeasibilityPolishingTermination = 5,
};  // enumerator order = the reference's (primal_dual_hybrid_gradient.h:70-83)
struct IterationCallbackInfo {
  IterationType iteration_type;
  TerminationCriteria termination_criteria;
  IterationStats iteration_stats;
  QuadraticProgramBoundNorms bound_norms;
};

// iteration_stats.h:89-102 (iteration_stats.cc:567-597): the entry of `stats` for one candidate point, if present.
inline std::optional<ConvergenceInformation> GetConvergenceInformation(const IterationStats& stats, int candidate_type) {
  for (const ConvergenceInformation& info : stats.convergence_information) if (info.candidate_type == candidate_type) return info;
  return std::nullopt;
}
inline std::optional<InfeasibilityInformation> GetInfeasibilityInformation(const IterationStats& stats, int candidate_type) {
  for (const InfeasibilityInformation& info : stats.infeasibility_information) if (info.candidate_type == candidate_type) return info;
  return std::nullopt;
}
inline std::optional<PointMetadata> GetPointMetadata(const IterationStats& stats, int point_type) {
  for (const PointMetadata& metadata : stats.point_metadata) if (metadata.point_type == point_type) return metadata;
  return std::nullopt;
}
// The generated-code spellings of the enum name tables (solve_log.pb.h).
inline const char* TerminationReason_Name(int reason) { return TerminationReasonName(reason); }
inline const char* PointType_Name(int point_type) { return PointTypeName(point_type); }

// Device placement; not part of the reference API (defaults reproduce it).
struct DeviceOptions {
  int device = -1;  // -1: the calling thread's current CUDA device
};

// Counters of the last solve on the calling thread, for benchmarks and tests.
struct SolveCounters {
  long long kernel_launches = 0;
  long long trial_steps = 0;
  double loop_seconds = 0.0;         // host wall time of Solver::Solve
  double device_loop_seconds = 0.0;  // CUDA-event time of the iteration loop
  double setup_seconds = 0.0;
  long long collectives = 0;         // NCCL all-reduces enqueued (0 on one GPU)
  int symmetric_buffers = 0;         // exchange buffers NCCL accepted as symmetric windows (0: plain all-reduce path)
  long long peer_exchanges = 0;      // exchange steps through peer memory (pdlp_peer.cuh) instead of ncclAllReduce
  double upload_seconds = 0.0;       // host -> device copies, CSR build, communicator set-up (before preprocessing)
  int interrupts_seen_on_device = 0; // times the decision kernel stopped the loop on the interrupt flag (one GPU; pdlp_step_state.h)
};
const SolveCounters& LastSolveCounters();

// primal_dual_hybrid_gradient.h:140-156.
SolverResult PrimalDualHybridGradient(QuadraticProgram qp, const PrimalDualHybridGradientParams& params,
                                      const std::atomic<bool>* interrupt_solve = nullptr,
                                      std::function<void(const IterationCallbackInfo&)> iteration_stats_callback = nullptr);
SolverResult PrimalDualHybridGradient(QuadraticProgram qp, const PrimalDualHybridGradientParams& params,
                                      std::optional<PrimalAndDualSolution> initial_solution,
                                      const std::atomic<bool>* interrupt_solve = nullptr,
                                      std::function<void(const IterationCallbackInfo&)> iteration_stats_callback = nullptr);

// Glop's basis statuses (ortools/lp_data/lp_types.h:201-241), with Glop's numbering, for the presolve hand-off: the
// reference turns a PDLP point into a glop::ProblemSolution with internal::ComputeStatuses before it asks Glop's
// preprocessor to recover the original solution (primal_dual_hybrid_gradient.cc:1295).  A caller's postsolve callback
// (pdlp_b200_solve_presolved) does the same with this restatement.
enum class VariableStatus : int8_t { BASIC = 0, FIXED_VALUE = 1, AT_LOWER_BOUND = 2, AT_UPPER_BOUND = 3, FREE = 4 };
enum class ConstraintStatus : int8_t { BASIC = 0, FIXED_VALUE = 1, AT_LOWER_BOUND = 2, AT_UPPER_BOUND = 3, FREE = 4 };
struct ProblemStatuses {  // the two status rows of glop::ProblemSolution
  std::vector<VariableStatus> variable_statuses;
  std::vector<ConstraintStatus> constraint_statuses;
};
// Glop's ProblemStatus (lp_types.h:105-160, Glop's numbering) and the table that turns a preprocessor's verdict into
// the TerminationReason handed to pdlp_b200_solve_presolved (GlopStatusToTerminationReason, pdhg.cc:1123-1143).
enum class GlopProblemStatus : int8_t {
  OPTIMAL = 0, PRIMAL_INFEASIBLE, DUAL_INFEASIBLE, INFEASIBLE_OR_UNBOUNDED, PRIMAL_UNBOUNDED, DUAL_UNBOUNDED, INIT, PRIMAL_FEASIBLE, DUAL_FEASIBLE,
  ABNORMAL, INVALID_PROBLEM, IMPRECISE
};
inline int GlopStatusToTerminationReason(GlopProblemStatus glop_status) {
  switch (glop_status) {
    case GlopProblemStatus::OPTIMAL: return TERMINATION_REASON_OPTIMAL;
    case GlopProblemStatus::INVALID_PROBLEM: return TERMINATION_REASON_INVALID_PROBLEM;
    case GlopProblemStatus::ABNORMAL:
    case GlopProblemStatus::IMPRECISE: return TERMINATION_REASON_NUMERICAL_ERROR;
    case GlopProblemStatus::PRIMAL_INFEASIBLE:
    case GlopProblemStatus::DUAL_INFEASIBLE:
    case GlopProblemStatus::INFEASIBLE_OR_UNBOUNDED:
    case GlopProblemStatus::DUAL_UNBOUNDED:
    case GlopProblemStatus::PRIMAL_UNBOUNDED: return TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE;
    default: return TERMINATION_REASON_OTHER;  // (the reference also logs a warning)
  }
}
namespace internal {
// primal_dual_hybrid_gradient.h:164, .cc:2835-2883.  Host code.
inline ProblemStatuses ComputeStatuses(const QuadraticProgram& qp, const PrimalAndDualSolution& solution) {
  ProblemStatuses out;
  out.constraint_statuses.resize(solution.dual_solution.size());
  out.variable_statuses.resize(solution.primal_solution.size());
  for (size_t i = 0; i < solution.dual_solution.size(); ++i) {
    if (qp.constraint_lower_bounds[i] == qp.constraint_upper_bounds[i]) out.constraint_statuses[i] = ConstraintStatus::FIXED_VALUE;
    else if (solution.dual_solution[i] > 0) out.constraint_statuses[i] = ConstraintStatus::AT_LOWER_BOUND;
    else if (solution.dual_solution[i] < 0) out.constraint_statuses[i] = ConstraintStatus::AT_UPPER_BOUND;
    else out.constraint_statuses[i] = ConstraintStatus::BASIC;
  }
  for (size_t j = 0; j < solution.primal_solution.size(); ++j) {
    const bool at_lb = solution.primal_solution[j] <= qp.variable_lower_bounds[j];
    const bool at_ub = solution.primal_solution[j] >= qp.variable_upper_bounds[j];
    out.variable_statuses[j] = at_lb ? (at_ub ? VariableStatus::FIXED_VALUE : VariableStatus::AT_LOWER_BOUND)
                                     : (at_ub ? VariableStatus::AT_UPPER_BOUND : VariableStatus::BASIC);
  }
  return out;
}
}  // namespace internal

}  // namespace pdlp_b200

// The reference's namespace (primal_dual_hybrid_gradient.h:28, quadratic_program.h:30).
namespace operations_research::pdlp {
using namespace ::pdlp_b200;  // NOLINT: qualified lookup of operations_research::pdlp::X finds pdlp_b200::X
}  // namespace operations_research::pdlp
#endif  // PDLP_B200_PRIMAL_DUAL_HYBRID_GRADIENT_H_
