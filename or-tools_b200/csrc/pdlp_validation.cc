// Range checks on PrimalDualHybridGradientParams; same checks, order and messages as
// ortools/pdlp/solvers_proto_validation.cc:45-298 (a failed nested check is annotated
// "; <field> invalid" the way util::StatusBuilder joins messages).
#include <cmath>
#include <cstdio>
#include <limits>

#include "pdlp_termination.h"

namespace pdlp_b200 {

namespace {

constexpr double kTinyDouble = 1.0e-50;
constexpr double kHugeDouble = 1.0e50;

std::string Num(double v) {  // absl::StrCat(double) prints like %g
  char b[64];
  std::snprintf(b, sizeof b, "%g", v);
  return b;
}

std::string CheckNonNegative(double v, const std::string& name) {
  if (std::isnan(v)) return name + " is NAN";
  if (v < 0) return name + " must be non-negative";
  return "";
}

#define PDLP_RETURN_IF_ERROR(expr)    \
  do {                                \
    std::string e__ = (expr);         \
    if (!e__.empty()) return e__;     \
  } while (0)

std::string ValidateTerminationCriteria(const TerminationCriteria& c) {
  const int norm = c.optimality_norm;
  if (norm != OPTIMALITY_NORM_L_INF && norm != OPTIMALITY_NORM_L2 && norm != OPTIMALITY_NORM_L_INF_COMPONENTWISE) return "invalid value for optimality_norm";
  if (c.detailed_optimality_criteria.has || c.simple_optimality_criteria.has) {
    if (c.eps_optimal_absolute.has) return "eps_optimal_absolute should not be set if detailed_optimality_criteria or simple_optimality_criteria is used";
    if (c.eps_optimal_relative.has) return "eps_optimal_relative should not be set if detailed_optimality_criteria or simple_optimality_criteria is used";
  }
  if (c.detailed_optimality_criteria.has) {
    const DetailedOptimalityCriteria& d = c.detailed_optimality_criteria.get();
    const std::string pre = "detailed_optimality_criteria.";
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_primal_residual_absolute, pre + "eps_optimal_primal_residual_absolute"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_primal_residual_relative, pre + "eps_optimal_primal_residual_relative"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_dual_residual_absolute, pre + "eps_optimal_dual_residual_absolute"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_dual_residual_relative, pre + "eps_optimal_dual_residual_relative"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_objective_gap_absolute, pre + "eps_optimal_objective_gap_absolute"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(d.eps_optimal_objective_gap_relative, pre + "eps_optimal_objective_gap_relative"));
  } else if (c.simple_optimality_criteria.has) {
    const SimpleOptimalityCriteria& s = c.simple_optimality_criteria.get();
    PDLP_RETURN_IF_ERROR(CheckNonNegative(s.eps_optimal_absolute, "simple_optimality_criteria.eps_optimal_absolute"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(s.eps_optimal_relative, "simple_optimality_criteria.eps_optimal_relative"));
  } else {
    PDLP_RETURN_IF_ERROR(CheckNonNegative(c.eps_optimal_absolute, "eps_optimal_absolute"));
    PDLP_RETURN_IF_ERROR(CheckNonNegative(c.eps_optimal_relative, "eps_optimal_relative"));
  }
  PDLP_RETURN_IF_ERROR(CheckNonNegative(c.eps_primal_infeasible, "eps_primal_infeasible"));
  PDLP_RETURN_IF_ERROR(CheckNonNegative(c.eps_dual_infeasible, "eps_dual_infeasible"));
  PDLP_RETURN_IF_ERROR(CheckNonNegative(c.time_sec_limit, "time_sec_limit"));
  if (c.iteration_limit < 0) return "iteration_limit must be non-negative";
  PDLP_RETURN_IF_ERROR(CheckNonNegative(c.kkt_matrix_pass_limit, "kkt_matrix_pass_limit"));
  return "";
}

std::string ValidateAdaptiveLinesearchParams(const AdaptiveLinesearchParams& p) {
  const double red = p.step_size_reduction_exponent, grow = p.step_size_growth_exponent;
  if (std::isnan(red)) return "step_size_reduction_exponent is NAN";
  if (red < 0.1 || red > 1.0) return "step_size_reduction_exponent must be between 0.1 and 1.0 inclusive";
  if (std::isnan(grow)) return "step_size_growth_exponent is NAN";
  if (grow < 0.1 || grow > 1.0) return "step_size_growth_exponent must be between 0.1 and 1.0 inclusive";
  return "";
}

std::string ValidateMalitskyPockParams(const MalitskyPockParams& p) {
  const double down = p.step_size_downscaling_factor, contr = p.linesearch_contraction_factor, interp = p.step_size_interpolation;
  if (std::isnan(down)) return "step_size_downscaling_factor is NAN";
  if (down <= kTinyDouble || down >= 1) return "step_size_downscaling_factor must be between " + Num(kTinyDouble) + " and 1 exclusive";
  if (std::isnan(contr)) return "linesearch_contraction_factor is NAN";
  if (contr <= 0 || contr >= 1) return "linesearch_contraction_factor must be between 0 and 1 exclusive";
  if (std::isnan(interp)) return "step_size_interpolation is NAN";
  if (interp < 0 || interp >= kHugeDouble) return "step_size_interpolation must be non-negative and less than " + Num(kHugeDouble);
  return "";
}

std::string ValidateBody(const PrimalDualHybridGradientParams& p) {
  std::string e = ValidateTerminationCriteria(p.termination_criteria.get());
  if (!e.empty()) return e + "; termination_criteria invalid";
  if (p.num_threads <= 0) return "num_threads must be positive";
  if (p.verbosity_level < 0) return "verbosity_level must be non-negative";
  if (p.log_interval_seconds < 0.0) return "log_interval_seconds must be non-negative";
  if (std::isnan(p.log_interval_seconds.get())) return "log_interval_seconds is NAN";
  if (p.major_iteration_frequency <= 0) return "major_iteration_frequency must be positive";
  if (p.termination_check_frequency <= 0) return "termination_check_frequency must be positive";
  const int rs = p.restart_strategy;
  if (rs != NO_RESTARTS && rs != EVERY_MAJOR_ITERATION && rs != ADAPTIVE_HEURISTIC && rs != ADAPTIVE_DISTANCE_BASED) return "invalid restart_strategy";
  const double smooth = p.primal_weight_update_smoothing;
  if (std::isnan(smooth)) return "primal_weight_update_smoothing is NAN";
  if (smooth < 0 || smooth > 1) return "primal_weight_update_smoothing must be between 0 and 1 inclusive";
  const double ipw = p.initial_primal_weight;
  if (std::isnan(ipw)) return "initial_primal_weight is NAN";
  if (p.initial_primal_weight.has && (ipw <= kTinyDouble || ipw >= kHugeDouble))
    return "initial_primal_weight must be between " + Num(kTinyDouble) + " and " + Num(kHugeDouble) + " if specified";
  if (p.l_inf_ruiz_iterations < 0) return "l_inf_ruiz_iterations must be non-negative";
  if (p.l_inf_ruiz_iterations > 100) return "l_inf_ruiz_iterations must be at most 100";
  const double suff = p.sufficient_reduction_for_restart, nec = p.necessary_reduction_for_restart;
  if (std::isnan(suff)) return "sufficient_reduction_for_restart is NAN";
  if (suff <= 0 || suff >= 1) return "sufficient_reduction_for_restart must be between 0 and 1 exclusive";
  if (std::isnan(nec)) return "necessary_reduction_for_restart is NAN";
  if (nec < suff || nec >= 1) return "necessary_reduction_for_restart must be in the interval [sufficient_reduction_for_restart, 1)";
  const int ls = p.linesearch_rule;
  if (ls != ADAPTIVE_LINESEARCH_RULE && ls != MALITSKY_POCK_LINESEARCH_RULE && ls != CONSTANT_STEP_SIZE_RULE) return "invalid linesearch_rule";
  e = ValidateAdaptiveLinesearchParams(p.adaptive_linesearch_parameters.get());
  if (!e.empty()) return e + "; adaptive_linesearch_parameters invalid";
  e = ValidateMalitskyPockParams(p.malitsky_pock_parameters.get());
  if (!e.empty()) return e + "; malitsky_pock_parameters invalid";
  const double iss = p.initial_step_size_scaling;
  if (std::isnan(iss)) return "initial_step_size_scaling is NAN";
  if (iss <= kTinyDouble || iss >= kHugeDouble) return "initial_step_size_scaling must be between " + Num(kTinyDouble) + " and " + Num(kHugeDouble);
  const double thr = p.infinite_constraint_bound_threshold;
  if (std::isnan(thr)) return "infinite_constraint_bound_threshold is NAN";
  if (thr <= 0.0) return "infinite_constraint_bound_threshold must be positive";
  const double tol = p.diagonal_qp_trust_region_solver_tolerance;
  if (std::isnan(tol)) return "diagonal_qp_trust_region_solver_tolerance is NAN";
  if (tol < 10 * std::numeric_limits<double>::epsilon())
    return "diagonal_qp_trust_region_solver_tolerance must be at least " + Num(10 * std::numeric_limits<double>::epsilon());
  if (p.use_feasibility_polishing && p.handle_some_primal_gradients_on_finite_bounds_as_residuals)
    return "use_feasibility_polishing requires !handle_some_primal_gradients_on_finite_bounds_as_residuals";
  if (p.use_feasibility_polishing && p.presolve_options->use_glop) return "use_feasibility_polishing and glop presolve can not be used together.";
  return "";
}

}  // namespace

std::string ValidatePrimalDualHybridGradientParams(const PrimalDualHybridGradientParams& p) {
  const std::string e = ValidateBody(p);
  return e.empty() ? e : "INVALID_ARGUMENT: " + e;
}

}  // namespace pdlp_b200
