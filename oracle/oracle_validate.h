// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// Parameter range checks.  Returns "" when valid, else the status text the reference
// would put in termination_string ("INVALID_ARGUMENT: <message>[; <context>]").
//   follows ortools/pdlp/solvers_proto_validation.cc:30-298,
//           ortools/base/status_builder.h (message "; " annotation join).
#ifndef PDLP_ORACLE_VALIDATE_H_
#define PDLP_ORACLE_VALIDATE_H_

#include "oracle_stats.h"

namespace oracle {
namespace detail {
inline std::string SixDigits(double v) { char b[64]; std::snprintf(b, sizeof b, "%g", v); return b; }  // absl::StrCat(double)
inline std::string NonNegative(double v, const std::string& name) {
  if (std::isnan(v)) return name + " is NAN";
  if (v < 0) return name + " must be non-negative";
  return "";
}
inline std::string ValidateTerminationCriteria(const TerminationCriteria& c) {
  if (c.optimality_norm != NORM_L_INF && c.optimality_norm != NORM_L2 && c.optimality_norm != NORM_L_INF_COMPONENTWISE) return "invalid value for optimality_norm";
  if (c.has_detailed || c.has_simple) {
    if (c.has_eps_optimal_absolute) return "eps_optimal_absolute should not be set if detailed_optimality_criteria or simple_optimality_criteria is used";
    if (c.has_eps_optimal_relative) return "eps_optimal_relative should not be set if detailed_optimality_criteria or simple_optimality_criteria is used";
  }
  std::string e;
#define ORACLE_RET(expr) do { e = (expr); if (!e.empty()) return e; } while (0)
  if (c.has_detailed) {
    const DetailedCriteria& d = c.detailed;
    ORACLE_RET(NonNegative(d.eps_optimal_primal_residual_absolute, "detailed_optimality_criteria.eps_optimal_primal_residual_absolute"));
    ORACLE_RET(NonNegative(d.eps_optimal_primal_residual_relative, "detailed_optimality_criteria.eps_optimal_primal_residual_relative"));
    ORACLE_RET(NonNegative(d.eps_optimal_dual_residual_absolute, "detailed_optimality_criteria.eps_optimal_dual_residual_absolute"));
    ORACLE_RET(NonNegative(d.eps_optimal_dual_residual_relative, "detailed_optimality_criteria.eps_optimal_dual_residual_relative"));
    ORACLE_RET(NonNegative(d.eps_optimal_objective_gap_absolute, "detailed_optimality_criteria.eps_optimal_objective_gap_absolute"));
    ORACLE_RET(NonNegative(d.eps_optimal_objective_gap_relative, "detailed_optimality_criteria.eps_optimal_objective_gap_relative"));
  } else if (c.has_simple) {
    ORACLE_RET(NonNegative(c.simple_eps_abs, "simple_optimality_criteria.eps_optimal_absolute"));
    ORACLE_RET(NonNegative(c.simple_eps_rel, "simple_optimality_criteria.eps_optimal_relative"));
  } else {
    ORACLE_RET(NonNegative(c.eps_optimal_absolute, "eps_optimal_absolute"));
    ORACLE_RET(NonNegative(c.eps_optimal_relative, "eps_optimal_relative"));
  }
  ORACLE_RET(NonNegative(c.eps_primal_infeasible, "eps_primal_infeasible"));
  ORACLE_RET(NonNegative(c.eps_dual_infeasible, "eps_dual_infeasible"));
  ORACLE_RET(NonNegative(c.time_sec_limit, "time_sec_limit"));
  if (c.iteration_limit < 0) return "iteration_limit must be non-negative";
  ORACLE_RET(NonNegative(c.kkt_matrix_pass_limit, "kkt_matrix_pass_limit"));
  return "";
}
inline std::string ValidateAdaptive(const Params& p) {
  if (std::isnan(p.step_size_reduction_exponent)) return "step_size_reduction_exponent is NAN";
  if (p.step_size_reduction_exponent < 0.1 || p.step_size_reduction_exponent > 1.0) return "step_size_reduction_exponent must be between 0.1 and 1.0 inclusive";
  if (std::isnan(p.step_size_growth_exponent)) return "step_size_growth_exponent is NAN";
  if (p.step_size_growth_exponent < 0.1 || p.step_size_growth_exponent > 1.0) return "step_size_growth_exponent must be between 0.1 and 1.0 inclusive";
  return "";
}
inline std::string ValidateMalitskyPock(const Params& p) {
  const double kTiny = 1.0e-50, kHuge = 1.0e50;
  if (std::isnan(p.step_size_downscaling_factor)) return "step_size_downscaling_factor is NAN";
  if (p.step_size_downscaling_factor <= kTiny || p.step_size_downscaling_factor >= 1) return "step_size_downscaling_factor must be between " + SixDigits(kTiny) + " and 1 exclusive";
  if (std::isnan(p.linesearch_contraction_factor)) return "linesearch_contraction_factor is NAN";
  if (p.linesearch_contraction_factor <= 0 || p.linesearch_contraction_factor >= 1) return "linesearch_contraction_factor must be between 0 and 1 exclusive";
  if (std::isnan(p.step_size_interpolation)) return "step_size_interpolation is NAN";
  if (p.step_size_interpolation < 0 || p.step_size_interpolation >= kHuge) return "step_size_interpolation must be non-negative and less than " + SixDigits(kHuge);
  return "";
}
inline std::string ValidateBody(const Params& p) {
  const double kTiny = 1.0e-50, kHuge = 1.0e50;
  std::string e = ValidateTerminationCriteria(p.termination_criteria);
  if (!e.empty()) return e + "; termination_criteria invalid";
  if (p.num_threads <= 0) return "num_threads must be positive";
  if (p.verbosity_level < 0) return "verbosity_level must be non-negative";
  if (p.log_interval_seconds < 0.0) return "log_interval_seconds must be non-negative";
  if (std::isnan(p.log_interval_seconds)) return "log_interval_seconds is NAN";
  if (p.major_iteration_frequency <= 0) return "major_iteration_frequency must be positive";
  if (p.termination_check_frequency <= 0) return "termination_check_frequency must be positive";
  if (p.restart_strategy != NO_RESTARTS && p.restart_strategy != EVERY_MAJOR_ITERATION && p.restart_strategy != ADAPTIVE_HEURISTIC && p.restart_strategy != ADAPTIVE_DISTANCE_BASED)
    return "invalid restart_strategy";
  if (std::isnan(p.primal_weight_update_smoothing)) return "primal_weight_update_smoothing is NAN";
  if (p.primal_weight_update_smoothing < 0 || p.primal_weight_update_smoothing > 1) return "primal_weight_update_smoothing must be between 0 and 1 inclusive";
  if (std::isnan(p.initial_primal_weight)) return "initial_primal_weight is NAN";
  if (p.has_initial_primal_weight && (p.initial_primal_weight <= kTiny || p.initial_primal_weight >= kHuge))
    return "initial_primal_weight must be between " + SixDigits(kTiny) + " and " + SixDigits(kHuge) + " if specified";
  if (p.l_inf_ruiz_iterations < 0) return "l_inf_ruiz_iterations must be non-negative";
  if (p.l_inf_ruiz_iterations > 100) return "l_inf_ruiz_iterations must be at most 100";
  if (std::isnan(p.sufficient_reduction_for_restart)) return "sufficient_reduction_for_restart is NAN";
  if (p.sufficient_reduction_for_restart <= 0 || p.sufficient_reduction_for_restart >= 1) return "sufficient_reduction_for_restart must be between 0 and 1 exclusive";
  if (std::isnan(p.necessary_reduction_for_restart)) return "necessary_reduction_for_restart is NAN";
  if (p.necessary_reduction_for_restart < p.sufficient_reduction_for_restart || p.necessary_reduction_for_restart >= 1)
    return "necessary_reduction_for_restart must be in the interval [sufficient_reduction_for_restart, 1)";
  if (p.linesearch_rule != ADAPTIVE_LINESEARCH_RULE && p.linesearch_rule != MALITSKY_POCK_LINESEARCH_RULE && p.linesearch_rule != CONSTANT_STEP_SIZE_RULE) return "invalid linesearch_rule";
  e = ValidateAdaptive(p);
  if (!e.empty()) return e + "; adaptive_linesearch_parameters invalid";
  e = ValidateMalitskyPock(p);
  if (!e.empty()) return e + "; malitsky_pock_parameters invalid";
  if (std::isnan(p.initial_step_size_scaling)) return "initial_step_size_scaling is NAN";
  if (p.initial_step_size_scaling <= kTiny || p.initial_step_size_scaling >= kHuge) return "initial_step_size_scaling must be between " + SixDigits(kTiny) + " and " + SixDigits(kHuge);
  if (std::isnan(p.infinite_constraint_bound_threshold)) return "infinite_constraint_bound_threshold is NAN";
  if (p.infinite_constraint_bound_threshold <= 0.0) return "infinite_constraint_bound_threshold must be positive";
  if (std::isnan(p.diagonal_qp_trust_region_solver_tolerance)) return "diagonal_qp_trust_region_solver_tolerance is NAN";
  if (p.diagonal_qp_trust_region_solver_tolerance < 10 * std::numeric_limits<double>::epsilon())
    return "diagonal_qp_trust_region_solver_tolerance must be at least " + SixDigits(10 * std::numeric_limits<double>::epsilon());
  if (p.use_feasibility_polishing && p.handle_some_primal_gradients_on_finite_bounds_as_residuals)
    return "use_feasibility_polishing requires !handle_some_primal_gradients_on_finite_bounds_as_residuals";
  if (p.use_feasibility_polishing && p.use_glop) return "use_feasibility_polishing and glop presolve can not be used together.";
  return "";
#undef ORACLE_RET
}
}  // namespace detail

inline std::string ValidateParams(const Params& p) {
  const std::string e = detail::ValidateBody(p);
  return e.empty() ? e : "INVALID_ARGUMENT: " + e;
}

}  // namespace oracle
#endif  // PDLP_ORACLE_VALIDATE_H_
