// Text logging of the solve under `verbosity_level` (solvers.proto:284-301), restated from
// ortools/pdlp/primal_dual_hybrid_gradient.cc:98-292 (the iteration table), :1220-1266 (problem
// statistics) and the call sites :982-986, :1053-1057, :1473-1501, :1617-1641, :2004, :2048-2060, :2200,
// :2457-2530.  Host code; lines go to stderr, where glog's LOG(INFO).NoPrefix() puts the reference's.
//   level 1  problem statistics before / after rescaling, termination summary
//   level 2  + one table line per termination check (a header every 15 lines), polishing messages
//   level 3  + kkt_pass column and the long convergence columns
//   level 4  + an iterate tag column, the current iterate's line after the average's, restart messages
#ifndef PDLP_B200_LOGGING_H_
#define PDLP_B200_LOGGING_H_

#include <cstdarg>
#include <cstdio>
#include <optional>
#include <string>
#include <vector>

#include "pdlp_b200/primal_dual_hybrid_gradient.h"

namespace pdlp_b200 {

inline std::string Format(const char* fmt, ...) {
  char buf[640];
  va_list ap;
  va_start(ap, fmt);
  std::vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  return buf;
}
inline void LogInfoWithoutPrefix(const std::string& message) {  // :184-186
  std::fputs((message + "\n").c_str(), stderr);
}

inline std::optional<ConvergenceInformation> GetConvergenceInformationFor(const IterationStats& stats, int candidate_type) {  // iteration_stats.cc:567-577
  for (const ConvergenceInformation& c : stats.convergence_information) {
    if (c.candidate_type == candidate_type) return c;
  }
  return std::nullopt;
}

// ConvergenceInformationString (:98-148) / ConvergenceInformationShortString (:150-182).
inline std::string ConvergenceColumns(bool long_form, const ConvergenceInformation& c, const RelativeConvergenceInformation& r, int norm) {
  double rel_p = r.relative_l2_primal_residual, rel_d = r.relative_l2_dual_residual, res_p = c.l2_primal_residual, res_d = c.l2_dual_residual;
  if (norm == OPTIMALITY_NORM_L_INF) {
    rel_p = r.relative_l_inf_primal_residual; rel_d = r.relative_l_inf_dual_residual; res_p = c.l_inf_primal_residual; res_d = c.l_inf_dual_residual;
  } else if (norm == OPTIMALITY_NORM_L_INF_COMPONENTWISE) {
    rel_p = c.l_inf_componentwise_primal_residual; rel_d = c.l_inf_componentwise_dual_residual; res_p = c.l_inf_primal_residual; res_d = c.l_inf_dual_residual;
  }
  if (long_form) {
    return Format("%#12.6g %#12.6g %#12.6g | %#12.6g %#12.6g %#12.6g | %#12.6g %#12.6g | %#12.6g %#12.6g", rel_p, rel_d, r.relative_optimality_gap, res_p, res_d,
                  c.primal_objective - c.dual_objective, c.primal_objective.get(), c.dual_objective.get(), c.l2_primal_variable.get(), c.l2_dual_variable.get());
  }
  return Format("%#10.4g %#10.4g %#10.4g | %#10.4g %#10.4g", rel_p, rel_d, r.relative_optimality_gap, c.primal_objective.get(), c.dual_objective.get());
}

// LogIterationStats (:193-266) as a string.
inline std::string IterationStatsLine(int verbosity_level, bool use_feasibility_polishing, IterationType iteration_type, const IterationStats& stats,
                                      const TerminationCriteria& criteria, const QuadraticProgramBoundNorms& bound_norms, int preferred_candidate) {
  const std::string iteration = verbosity_level >= 3 ? Format("%6d %8.1f %6.1f", static_cast<int>(stats.iteration_number), stats.cumulative_kkt_matrix_passes.get(),
                                                              stats.cumulative_time_sec.get())
                                                     : Format("%6d %6.1f", static_cast<int>(stats.iteration_number), stats.cumulative_time_sec.get());
  std::optional<ConvergenceInformation> info = GetConvergenceInformationFor(stats, preferred_candidate);
  if (!info.has_value() && !stats.convergence_information.empty()) info = stats.convergence_information[0];
  const char* phase = "";
  if (use_feasibility_polishing) {
    switch (iteration_type) {
      case IterationType::kNormal: phase = "n "; break;
      case IterationType::kPrimalFeasibility: phase = "p "; break;
      case IterationType::kDualFeasibility: phase = "d "; break;
      default: phase = "t "; break;
    }
  }
  if (!info.has_value()) return std::string(phase) + (verbosity_level >= 4 ? "? " : "") + iteration;
  const char* iterate = "";
  if (verbosity_level >= 4) {
    switch (static_cast<int>(info->candidate_type)) {
      case POINT_TYPE_CURRENT_ITERATE: iterate = "C "; break;
      case POINT_TYPE_AVERAGE_ITERATE: iterate = "A "; break;
      case POINT_TYPE_ITERATE_DIFFERENCE: iterate = "D "; break;
      case POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION: iterate = "F "; break;
      default: iterate = "? "; break;
    }
  }
  const RelativeConvergenceInformation rel = ComputeRelativeResiduals(EffectiveOptimalityCriteria(criteria), *info, bound_norms);
  return std::string(phase) + iterate + iteration + " | " + ConvergenceColumns(verbosity_level >= 3, *info, rel, criteria.optimality_norm);
}

// LogIterationStatsHeader (:268-287) as a string.
inline std::string IterationStatsHeader(int verbosity_level, bool use_feasibility_polishing) {
  const std::string work = verbosity_level >= 3 ? Format("%6s %8s %6s", "iter#", "kkt_pass", "time") : Format("%6s %6s", "iter#", "time");
  const std::string conv = verbosity_level >= 3
                               ? Format("%12s %12s %12s | %12s %12s %12s | %12s %12s | %12s %12s", "rel_prim_res", "rel_dual_res", "rel_gap", "prim_resid", "dual_resid",
                                        "obj_gap", "prim_obj", "dual_obj", "prim_var_l2", "dual_var_l2")
                               : Format("%10s %10s %10s | %10s %10s", "rel_p_res", "rel_d_res", "rel_gap", "prim_obj", "dual_obj");
  return std::string(use_feasibility_polishing ? "f " : "") + (verbosity_level >= 4 ? "I " : "") + work + " | " + conv;
}

// LogQuadraticProgramStats (:1220-1266) as lines.
inline std::vector<std::string> QuadraticProgramStatsLines(const QuadraticProgramStats& s, bool has_nonzeros, bool is_lp) {
  std::vector<std::string> out;
  out.push_back(Format("There are %lld variables, %lld constraints, and %lld constraint matrix nonzeros.", static_cast<long long>(s.num_variables),
                       static_cast<long long>(s.num_constraints), static_cast<long long>(s.constraint_matrix_num_nonzeros)));
  if (has_nonzeros) {
    out.push_back(Format("Absolute values of nonzero constraint matrix elements: largest=%f, smallest=%f, avg=%f", s.constraint_matrix_abs_max.get(),
                         s.constraint_matrix_abs_min.get(), s.constraint_matrix_abs_avg.get()));
    out.push_back(Format("Constraint matrix, infinity norm: max(row & col)=%f, min_col=%f, min_row=%f", s.constraint_matrix_abs_max.get(),
                         s.constraint_matrix_col_min_l_inf_norm.get(), s.constraint_matrix_row_min_l_inf_norm.get()));
    out.push_back(Format("Constraint bounds statistics (max absolute value per row): largest=%f, smallest=%f, avg=%f, l2_norm=%f", s.combined_bounds_max.get(),
                         s.combined_bounds_min.get(), s.combined_bounds_avg.get(), s.combined_bounds_l2_norm.get()));
  }
  if (!is_lp) {
    out.push_back(Format("There are %lld nonzero diagonal coefficients in the objective matrix.", static_cast<long long>(s.objective_matrix_num_nonzeros)));
    out.push_back(Format("Absolute values of nonzero objective matrix elements: largest=%f, smallest=%f, avg=%f", s.objective_matrix_abs_max.get(),
                         s.objective_matrix_abs_min.get(), s.objective_matrix_abs_avg.get()));
  }
  out.push_back(Format("Absolute values of objective vector elements: largest=%f, smallest=%f, avg=%f, l2_norm=%f", s.objective_vector_abs_max.get(),
                       s.objective_vector_abs_min.get(), s.objective_vector_abs_avg.get(), s.objective_vector_l2_norm.get()));
  out.push_back(Format("Gaps between variable upper and lower bounds: #finite=%lld of %lld, largest=%f, smallest=%f, avg=%f",
                       static_cast<long long>(s.variable_bound_gaps_num_finite), static_cast<long long>(s.num_variables), s.variable_bound_gaps_max.get(),
                       s.variable_bound_gaps_min.get(), s.variable_bound_gaps_avg.get()));
  return out;
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_LOGGING_H_
