// Host-side scalar logic on the schema records: termination predicates
// (ortools/pdlp/termination.h:25-130, termination.cc:26-271) and parameter range checks
// (ortools/pdlp/solvers_proto_validation.cc:45-298).  Pure functions of a handful of
// doubles; nothing here touches vectors.
#ifndef PDLP_B200_TERMINATION_H_
#define PDLP_B200_TERMINATION_H_

#include <atomic>
#include <cmath>
#include <cstdio>
#include <limits>
#include <optional>
#include <string>

#include "pdlp_schema.h"

namespace pdlp_b200 {

struct QuadraticProgramBoundNorms {  // termination.h:36-41
  double l2_norm_primal_linear_objective = 0, l2_norm_constraint_bounds = 0;
  double l_inf_norm_primal_linear_objective = 0, l_inf_norm_constraint_bounds = 0;
};
struct TerminationReasonAndPointType { int reason; int type; };
// The reference's `const std::atomic<bool>* interrupt_solve` (pdhg.h:140-156); the C ABI
// hands the same thing over as a `volatile int`.
struct InterruptFlag {
  const std::atomic<bool>* atomic_flag = nullptr;
  const volatile int* int_flag = nullptr;
  int agreed = -1;  // sharded solves: the value all ranks agreed on at this host event (>= 0 overrides the flags)
  bool IsSet() const {
    if (agreed >= 0) return agreed != 0;
    return (atomic_flag != nullptr && atomic_flag->load()) || (int_flag != nullptr && *int_flag != 0);
  }
};
struct RelativeConvergenceInformation {
  double relative_l_inf_primal_residual = 0, relative_l2_primal_residual = 0;
  double relative_l_inf_dual_residual = 0, relative_l2_dual_residual = 0, relative_optimality_gap = 0;
};

// termination.cc:126-159: detailed criteria in effect, whichever way they were given.
inline DetailedOptimalityCriteria EffectiveOptimalityCriteria(const TerminationCriteria& tc) {
  if (tc.detailed_optimality_criteria.has) return tc.detailed_optimality_criteria.get();
  double a = tc.eps_optimal_absolute, r = tc.eps_optimal_relative;
  if (tc.simple_optimality_criteria.has) {
    a = tc.simple_optimality_criteria->eps_optimal_absolute;
    r = tc.simple_optimality_criteria->eps_optimal_relative;
  }
  DetailedOptimalityCriteria d;
  d.eps_optimal_primal_residual_absolute = a; d.eps_optimal_primal_residual_relative = r;
  d.eps_optimal_dual_residual_absolute = a; d.eps_optimal_dual_residual_relative = r;
  d.eps_optimal_objective_gap_absolute = a; d.eps_optimal_objective_gap_relative = r;
  return d;
}

inline bool ObjectiveGapMet(const DetailedOptimalityCriteria& oc, const ConvergenceInformation& s) {  // termination.cc:26-41
  const double ea = oc.eps_optimal_objective_gap_absolute, er = oc.eps_optimal_objective_gap_relative;
  if (std::isinf(ea) || std::isinf(er)) return true;
  const double abs_obj = std::abs(s.primal_objective.get()) + std::abs(s.dual_objective.get());
  const double gap = std::abs(s.primal_objective.get() - s.dual_objective.get());
  return std::isfinite(abs_obj) && gap <= ea + er * abs_obj;
}

inline bool OptimalityCriteriaMet(const DetailedOptimalityCriteria& oc, const ConvergenceInformation& s, int norm, const QuadraticProgramBoundNorms& bn) {  // :43-97
  double perr = 0, pbase = 0, derr = 0, dbase = 0;
  double pabs = oc.eps_optimal_primal_residual_absolute, dabs = oc.eps_optimal_dual_residual_absolute;
  switch (norm) {
    case OPTIMALITY_NORM_L_INF:
      perr = s.l_inf_primal_residual; pbase = bn.l_inf_norm_constraint_bounds;
      derr = s.l_inf_dual_residual; dbase = bn.l_inf_norm_primal_linear_objective; break;
    case OPTIMALITY_NORM_L2:
      perr = s.l2_primal_residual; pbase = bn.l2_norm_constraint_bounds;
      derr = s.l2_dual_residual; dbase = bn.l2_norm_primal_linear_objective; break;
    case OPTIMALITY_NORM_L_INF_COMPONENTWISE:
      perr = s.l_inf_componentwise_primal_residual; pbase = 1.0; pabs = 0.0;
      derr = s.l_inf_componentwise_dual_residual; dbase = 1.0; dabs = 0.0; break;
    default:
      std::fprintf(stderr, "pdlp_b200: invalid optimality_norm %d\n", norm);
      std::abort();
  }
  const double prel = oc.eps_optimal_primal_residual_relative, drel = oc.eps_optimal_dual_residual_relative;
  const bool p_ok = std::isinf(oc.eps_optimal_primal_residual_absolute.get()) || std::isinf(prel) || perr <= pabs + prel * pbase;
  const bool d_ok = std::isinf(oc.eps_optimal_dual_residual_absolute.get()) || std::isinf(drel) || derr <= dabs + drel * dbase;
  return p_ok && d_ok && ObjectiveGapMet(oc, s);
}

inline bool PrimalInfeasibilityCriteriaMet(double eps, const InfeasibilityInformation& s) {  // :104-109
  if (s.dual_ray_objective <= 0.0) return false;
  return s.max_dual_ray_infeasibility / s.dual_ray_objective <= eps;
}
inline bool DualInfeasibilityCriteriaMet(double eps, const InfeasibilityInformation& s) {  // :113-122
  if (s.primal_ray_linear_objective >= 0.0) return false;
  return (s.max_primal_ray_infeasibility / -s.primal_ray_linear_objective <= eps) &&
         (s.primal_ray_quadratic_norm / -s.primal_ray_linear_objective <= eps);
}

// termination.cc:161-184.  Work limits and the user's interrupt flag.
inline std::optional<TerminationReasonAndPointType> CheckSimpleTerminationCriteria(const TerminationCriteria& c, const IterationStats& s,
                                                                                  const InterruptFlag* interrupt = nullptr) {
  if (s.iteration_number >= c.iteration_limit) return TerminationReasonAndPointType{TERMINATION_REASON_ITERATION_LIMIT, POINT_TYPE_NONE};
  if (s.cumulative_kkt_matrix_passes >= c.kkt_matrix_pass_limit) return TerminationReasonAndPointType{TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT, POINT_TYPE_NONE};
  if (s.cumulative_time_sec >= c.time_sec_limit) return TerminationReasonAndPointType{TERMINATION_REASON_TIME_LIMIT, POINT_TYPE_NONE};
  if (interrupt != nullptr && interrupt->IsSet()) return TerminationReasonAndPointType{TERMINATION_REASON_INTERRUPTED_BY_USER, POINT_TYPE_NONE};
  return std::nullopt;
}

// termination.cc:186-219.  The first candidate that passes wins (current before average).
inline std::optional<TerminationReasonAndPointType> CheckIterateTerminationCriteria(const TerminationCriteria& c, const IterationStats& s,
                                                                                   const QuadraticProgramBoundNorms& bn,
                                                                                   bool force_numerical_termination = false) {
  const DetailedOptimalityCriteria oc = EffectiveOptimalityCriteria(c);
  for (const ConvergenceInformation& ci : s.convergence_information)
    if (OptimalityCriteriaMet(oc, ci, c.optimality_norm, bn)) return TerminationReasonAndPointType{TERMINATION_REASON_OPTIMAL, ci.candidate_type};
  for (const InfeasibilityInformation& ii : s.infeasibility_information) {
    if (PrimalInfeasibilityCriteriaMet(c.eps_primal_infeasible, ii)) return TerminationReasonAndPointType{TERMINATION_REASON_PRIMAL_INFEASIBLE, ii.candidate_type};
    if (DualInfeasibilityCriteriaMet(c.eps_dual_infeasible, ii)) return TerminationReasonAndPointType{TERMINATION_REASON_DUAL_INFEASIBLE, ii.candidate_type};
  }
  if (force_numerical_termination) return TerminationReasonAndPointType{TERMINATION_REASON_NUMERICAL_ERROR, POINT_TYPE_NONE};
  return std::nullopt;
}

inline QuadraticProgramBoundNorms BoundNormsFromProblemStats(const QuadraticProgramStats& s) {  // :221-228
  return {s.objective_vector_l2_norm, s.combined_bounds_l2_norm, s.objective_vector_abs_max, s.combined_bounds_max};
}
inline double EpsilonRatio(double eps_abs, double eps_rel) { return (eps_abs == eps_rel) ? 1.0 : eps_abs / eps_rel; }  // :230-237

inline RelativeConvergenceInformation ComputeRelativeResiduals(const DetailedOptimalityCriteria& oc, const ConvergenceInformation& s,
                                                               const QuadraticProgramBoundNorms& bn) {  // :239-271
  const double rp = EpsilonRatio(oc.eps_optimal_primal_residual_absolute, oc.eps_optimal_primal_residual_relative);
  const double rd = EpsilonRatio(oc.eps_optimal_dual_residual_absolute, oc.eps_optimal_dual_residual_relative);
  const double rg = EpsilonRatio(oc.eps_optimal_objective_gap_absolute, oc.eps_optimal_objective_gap_relative);
  RelativeConvergenceInformation r;
  r.relative_l_inf_primal_residual = s.l_inf_primal_residual / (rp + bn.l_inf_norm_constraint_bounds);
  r.relative_l2_primal_residual = s.l2_primal_residual / (rp + bn.l2_norm_constraint_bounds);
  r.relative_l_inf_dual_residual = s.l_inf_dual_residual / (rd + bn.l_inf_norm_primal_linear_objective);
  r.relative_l2_dual_residual = s.l2_dual_residual / (rd + bn.l2_norm_primal_linear_objective);
  const double abs_obj = std::abs(s.primal_objective.get()) + std::abs(s.dual_objective.get());
  r.relative_optimality_gap = (s.primal_objective - s.dual_objective) / (rg + abs_obj);
  return r;
}

// Returns "" when valid, else "INVALID_ARGUMENT: <message>" as the reference's
// absl::Status would print (solvers_proto_validation.cc:45-298).
std::string ValidatePrimalDualHybridGradientParams(const PrimalDualHybridGradientParams& p);

}  // namespace pdlp_b200
#endif  // PDLP_B200_TERMINATION_H_
