// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// Parameters (flat mirror of PrimalDualHybridGradientParams), solve-log records,
// KKT residual / objective statistics, termination predicates, parameter validation.
//   follows ortools/pdlp/solvers.proto:20-481, ortools/pdlp/solve_log.proto:24-450,
//           ortools/pdlp/iteration_stats.cc:66-610, ortools/pdlp/termination.cc:26-271,
//           ortools/pdlp/solvers_proto_validation.cc:45-298.
#ifndef PDLP_ORACLE_STATS_H_
#define PDLP_ORACLE_STATS_H_

#include <sstream>

#include "oracle_qp.h"

namespace oracle {

// Enum numbers as in solvers.proto / solve_log.proto.
enum { NORM_UNSPECIFIED = 0, NORM_L_INF = 1, NORM_L2 = 2, NORM_L_INF_COMPONENTWISE = 3 };
enum { RESTART_STRATEGY_UNSPECIFIED = 0, NO_RESTARTS = 1, EVERY_MAJOR_ITERATION = 2, ADAPTIVE_HEURISTIC = 3, ADAPTIVE_DISTANCE_BASED = 4 };
enum { LINESEARCH_UNSPECIFIED = 0, ADAPTIVE_LINESEARCH_RULE = 1, MALITSKY_POCK_LINESEARCH_RULE = 2, CONSTANT_STEP_SIZE_RULE = 3 };
enum { RESTART_CHOICE_UNSPECIFIED = 0, RESTART_CHOICE_NO_RESTART = 1, RESTART_CHOICE_WEIGHTED_AVERAGE_RESET = 2, RESTART_CHOICE_RESTART_TO_AVERAGE = 3 };
enum { POINT_TYPE_UNSPECIFIED = 0, POINT_TYPE_CURRENT_ITERATE = 1, POINT_TYPE_ITERATE_DIFFERENCE = 2, POINT_TYPE_AVERAGE_ITERATE = 3, POINT_TYPE_NONE = 4,
       POINT_TYPE_PRESOLVER_SOLUTION = 5, POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION = 6 };
enum { TERMINATION_REASON_UNSPECIFIED = 0, TERMINATION_REASON_OPTIMAL = 1, TERMINATION_REASON_PRIMAL_INFEASIBLE = 2, TERMINATION_REASON_DUAL_INFEASIBLE = 3,
       TERMINATION_REASON_TIME_LIMIT = 4, TERMINATION_REASON_ITERATION_LIMIT = 5, TERMINATION_REASON_NUMERICAL_ERROR = 6, TERMINATION_REASON_OTHER = 7,
       TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT = 8, TERMINATION_REASON_INVALID_PROBLEM = 9, TERMINATION_REASON_INVALID_PARAMETER = 10,
       TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE = 11, TERMINATION_REASON_INTERRUPTED_BY_USER = 12, TERMINATION_REASON_INVALID_INITIAL_SOLUTION = 13 };
enum { POLISHING_PHASE_TYPE_UNSPECIFIED = 0, POLISHING_PHASE_TYPE_PRIMAL_FEASIBILITY = 1, POLISHING_PHASE_TYPE_DUAL_FEASIBILITY = 2 };

struct DetailedCriteria {  // solvers.proto TerminationCriteria.DetailedOptimalityCriteria
  double eps_optimal_primal_residual_absolute = 1e-6, eps_optimal_primal_residual_relative = 1e-6;
  double eps_optimal_dual_residual_absolute = 1e-6, eps_optimal_dual_residual_relative = 1e-6;
  double eps_optimal_objective_gap_absolute = 1e-6, eps_optimal_objective_gap_relative = 1e-6;
};

struct TerminationCriteria {
  int optimality_norm = NORM_L2;
  bool has_simple = false; double simple_eps_abs = 1e-6, simple_eps_rel = 1e-6;
  bool has_detailed = false; DetailedCriteria detailed;
  bool has_eps_optimal_absolute = false, has_eps_optimal_relative = false;
  double eps_optimal_absolute = 1e-6, eps_optimal_relative = 1e-6;
  double eps_primal_infeasible = 1e-8, eps_dual_infeasible = 1e-8;
  double time_sec_limit = kInf;
  int32_t iteration_limit = 2147483647;
  double kkt_matrix_pass_limit = kInf;
};

struct Params {
  TerminationCriteria termination_criteria;
  int num_threads = 1, num_shards = 0;
  bool record_iteration_stats = false;
  int verbosity_level = 0;
  double log_interval_seconds = 0.0;
  int major_iteration_frequency = 64, termination_check_frequency = 64;
  int restart_strategy = ADAPTIVE_HEURISTIC;
  double primal_weight_update_smoothing = 0.5;
  bool has_initial_primal_weight = false; double initial_primal_weight = 0.0;
  bool use_glop = false;
  int l_inf_ruiz_iterations = 5;
  bool l2_norm_rescaling = true;
  double sufficient_reduction_for_restart = 0.1, necessary_reduction_for_restart = 0.9;
  int linesearch_rule = ADAPTIVE_LINESEARCH_RULE;
  double step_size_reduction_exponent = 0.3, step_size_growth_exponent = 0.6;  // adaptive_linesearch_parameters
  double step_size_downscaling_factor = 0.7, linesearch_contraction_factor = 0.99, step_size_interpolation = 1.0;  // malitsky_pock_parameters
  double initial_step_size_scaling = 1.0;
  std::vector<int> random_projection_seeds;
  double infinite_constraint_bound_threshold = kInf;
  bool handle_some_primal_gradients_on_finite_bounds_as_residuals = true;
  bool use_diagonal_qp_trust_region_solver = false;
  double diagonal_qp_trust_region_solver_tolerance = 1e-8;
  bool use_feasibility_polishing = false;
};

// ---- solve_log.proto records ----
struct ConvergenceInformation {
  int candidate_type = 0;
  double primal_objective = 0, dual_objective = 0, corrected_dual_objective = 0;
  double l_inf_primal_residual = 0, l2_primal_residual = 0, l_inf_componentwise_primal_residual = 0;
  double l_inf_dual_residual = 0, l2_dual_residual = 0, l_inf_componentwise_dual_residual = 0;
  double l_inf_primal_variable = 0, l2_primal_variable = 0, l_inf_dual_variable = 0, l2_dual_variable = 0;
};
struct InfeasibilityInformation {
  double max_primal_ray_infeasibility = 0, primal_ray_linear_objective = 0, primal_ray_quadratic_norm = 0;
  double max_dual_ray_infeasibility = 0, dual_ray_objective = 0;
  int candidate_type = 0;
};
struct PointMetadata {
  int point_type = 0;
  Vec random_primal_projections, random_dual_projections;
  bool has_active = false;
  int64_t active_primal_variable_count = 0, active_dual_variable_count = 0, active_primal_variable_change = 0, active_dual_variable_change = 0;
};
struct IterationStats {
  int iteration_number = 0;
  std::vector<ConvergenceInformation> convergence_information;
  std::vector<InfeasibilityInformation> infeasibility_information;
  std::vector<PointMetadata> point_metadata;
  double cumulative_kkt_matrix_passes = 0;
  int cumulative_rejected_steps = 0;
  double cumulative_time_sec = 0;
  int restart_used = 0;
  double step_size = 0, primal_weight = 0;
};
struct SolveLog;
struct FeasibilityPolishingDetails {
  int polishing_phase_type = 0, main_iteration_count = 0;
  DetailedCriteria detailed_criteria;  // params.termination_criteria.detailed_optimality_criteria of the phase
  int32_t iteration_limit = 0;         // params.termination_criteria.iteration_limit of the phase
  int termination_reason = 0, iteration_count = 0;
  double solve_time_sec = 0;
  IterationStats solution_stats;
  int solution_type = 0;
  std::vector<IterationStats> iteration_stats;
};
struct SolveLog {
  std::string instance_name;
  int termination_reason = 0;
  std::string termination_string;
  int iteration_count = 0;
  double preprocessing_time_sec = 0, solve_time_sec = 0;
  IterationStats solution_stats;
  int solution_type = 0;
  std::vector<IterationStats> iteration_stats;
  bool has_original_stats = false, has_preprocessed_stats = false;
  QpStats original_problem_stats, preprocessed_problem_stats;
  std::vector<FeasibilityPolishingDetails> feasibility_polishing_details;
};

// ---- iteration_stats.cc ----
struct ResidualNorms { double objective_correction, objective_full_correction, l_inf_residual, l_2_residual, l_inf_componentwise_residual; };

// iteration_stats.cc:66-134.
inline ResidualNorms PrimalResidualNorms(const ShardedQp& sq, const Vec& row_scale, const Vec& x, double componentwise_offset, bool homogeneous = false) {
  const Qp& qp = sq.qp();
  const Vec kx = TransposedMatVec(sq.Kt(), x, sq.KtSharder());
  const int ns = sq.DualSharder().NumShards();
  Vec linf(ns, 0.0), sumsq(ns, 0.0), lcw(ns, 0.0);
  sq.DualSharder().ForEach([&](int s, int64_t b, int64_t e) {
    double a_inf = 0, a_sq = 0, a_cw = 0;
    for (int64_t i = b; i < e; ++i) {
      const double ub = (homogeneous && std::isfinite(qp.cub[i])) ? 0.0 : qp.cub[i];
      const double lb = (homogeneous && std::isfinite(qp.clb[i])) ? 0.0 : qp.clb[i];
      double scaled_res = 0.0, res_bound = 0.0;
      if (kx[i] > ub) { scaled_res = kx[i] - ub; res_bound = ub; }
      else if (kx[i] < lb) { scaled_res = lb - kx[i]; res_bound = lb; }
      const double res = scaled_res / row_scale[i];
      a_inf = std::max(a_inf, res);
      a_sq += res * res;
      if (res > 0.0) a_cw = std::max(a_cw, res / (componentwise_offset + std::abs(res_bound / row_scale[i])));
    }
    linf[s] = a_inf; sumsq[s] = a_sq; lcw[s] = a_cw;
  });
  double t_inf = 0, t_sq = 0, t_cw = 0;
  for (int s = 0; s < ns; ++s) { t_inf = std::max(t_inf, linf[s]); t_sq += sumsq[s]; t_cw = std::max(t_cw, lcw[s]); }
  return {0.0, 0.0, t_inf, std::sqrt(t_sq), t_cw};
}

// iteration_stats.cc:140-151.
inline bool HandlePrimalGradientTermAsReducedCost(const Params& p, double g, double x, double lb, double ub) {
  if (g == 0.0) return true;
  const double active = g > 0.0 ? lb : ub;
  if (p.handle_some_primal_gradients_on_finite_bounds_as_residuals) return std::abs(x - active) <= std::abs(x);
  return std::isfinite(active);
}

// iteration_stats.cc:163-236.
inline ResidualNorms DualResidualNorms(const Params& p, const ShardedQp& sq, const Vec& col_scale, const Vec& x, const Vec& grad, double componentwise_offset) {
  const Qp& qp = sq.qp();
  const int ns = sq.PrimalSharder().NumShards();
  Vec corr(ns, 0.0), full(ns, 0.0), linf(ns, 0.0), sumsq(ns, 0.0), lcw(ns, 0.0);
  sq.PrimalSharder().ForEach([&](int s, int64_t b, int64_t e) {
    double a_corr = 0, a_full = 0, a_inf = 0, a_sq = 0, a_cw = 0;
    for (int64_t j = b; j < e; ++j) {
      if (grad[j] == 0.0) continue;
      const double bound = grad[j] > 0.0 ? qp.vlb[j] : qp.vub[j];
      a_full += bound * grad[j];
      if (HandlePrimalGradientTermAsReducedCost(p, grad[j], x[j], qp.vlb[j], qp.vub[j])) {
        a_corr += bound * grad[j];
      } else {
        const double res = std::abs(grad[j]) / col_scale[j];
        a_inf = std::max(a_inf, res);
        a_sq += res * res;
        if (res > 0.0) a_cw = std::max(a_cw, res / (componentwise_offset + std::abs(qp.c[j] / col_scale[j])));
      }
    }
    corr[s] = a_corr; full[s] = a_full; linf[s] = a_inf; sumsq[s] = a_sq; lcw[s] = a_cw;
  });
  double t_corr = 0, t_full = 0, t_inf = 0, t_sq = 0, t_cw = 0;
  for (int s = 0; s < ns; ++s) { t_corr += corr[s]; t_full += full[s]; t_inf = std::max(t_inf, linf[s]); t_sq += sumsq[s]; t_cw = std::max(t_cw, lcw[s]); }
  return {t_corr, t_full, t_inf, std::sqrt(t_sq), t_cw};
}

inline Vec ObjectiveProduct(const ShardedQp& sq, const Vec& x) {  // :239-253
  Vec r(x.size());
  if (sq.qp().IsLp()) { SetZero(sq.PrimalSharder(), r); return r; }
  const Vec& q = *sq.qp().q;
  sq.PrimalSharder().ForEach([&](int, int64_t b, int64_t e) { for (int64_t j = b; j < e; ++j) r[j] = q[j] * x[j]; });
  return r;
}
inline double QuadraticObjective(const ShardedQp& sq, const Vec& x, const Vec& qx) { return 0.5 * Dot(qx, x, sq.PrimalSharder()); }  // :256-263

// Qx + c - K^T y, or -K^T y with a zero primal objective (iteration_stats.cc:268-289).
inline Vec PrimalGradientFromObjectiveProduct(const ShardedQp& sq, const Vec& y, Vec qx, bool zero_objective = false) {
  const Qp& qp = sq.qp();
  sq.KSharder().ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j) {
      double kty = 0.0;
      for (int64_t k = qp.K.ptr[j]; k < qp.K.ptr[j + 1]; ++k) kty += qp.K.val[k] * y[qp.K.idx[k]];
      if (zero_objective) qx[j] = -kty;
      else qx[j] += qp.c[j] - kty;
    }
  });
  return qx;
}

inline double DualObjectiveBoundsTerm(const ShardedQp& sq, const Vec& y) {  // :294-316
  const Qp& qp = sq.qp();
  return sq.DualSharder().Sum([&](int, int64_t b, int64_t e) {
    double s = 0.0;
    for (int64_t i = b; i < e; ++i) {
      if (y[i] > 0.0) s += qp.clb[i] * y[i];
      else if (y[i] < 0.0) s += qp.cub[i] * y[i];
    }
    return s;
  });
}

// iteration_stats.cc:349-419.
inline ConvergenceInformation ComputeConvergenceInformation(const Params& p, const ShardedQp& sq, const Vec& col_scale, const Vec& row_scale, const Vec& x,
                                                            const Vec& y, double cw_primal_offset, double cw_dual_offset, int candidate_type) {
  const Qp& qp = sq.qp();
  ConvergenceInformation r;
  const ResidualNorms pr = PrimalResidualNorms(sq, row_scale, x, cw_primal_offset);
  r.l_inf_primal_residual = pr.l_inf_residual;
  r.l2_primal_residual = pr.l_2_residual;
  r.l_inf_componentwise_primal_residual = pr.l_inf_componentwise_residual;
  r.l_inf_primal_variable = ScaledLInfNorm(x, col_scale, sq.PrimalSharder());
  r.l2_primal_variable = ScaledNorm(x, col_scale, sq.PrimalSharder());
  r.l_inf_dual_variable = ScaledLInfNorm(y, row_scale, sq.DualSharder());
  r.l2_dual_variable = ScaledNorm(y, row_scale, sq.DualSharder());
  Vec qx = ObjectiveProduct(sq, x);
  const double quad = QuadraticObjective(sq, x, qx);
  const Vec grad = PrimalGradientFromObjectiveProduct(sq, y, std::move(qx));
  r.primal_objective = qp.ApplyObjectiveScalingAndOffset(quad + Dot(qp.c, x, sq.PrimalSharder()));
  const double dual_piece = -quad + DualObjectiveBoundsTerm(sq, y);
  const ResidualNorms dr = DualResidualNorms(p, sq, col_scale, x, grad, cw_dual_offset);
  r.dual_objective = qp.ApplyObjectiveScalingAndOffset(dual_piece + dr.objective_correction);
  r.corrected_dual_objective = qp.ApplyObjectiveScalingAndOffset(dual_piece + dr.objective_full_correction);
  r.l_inf_dual_residual = dr.l_inf_residual;
  r.l2_dual_residual = dr.l_2_residual;
  r.l_inf_componentwise_dual_residual = dr.l_inf_componentwise_residual;
  r.candidate_type = candidate_type;
  return r;
}

// iteration_stats.cc:421-518.
inline InfeasibilityInformation ComputeInfeasibilityInformation(const Params& p, const ShardedQp& sq, const Vec& col_scale, const Vec& row_scale,
                                                                const Vec& x_ray, const Vec& y_ray, int candidate_type) {
  const Qp& qp = sq.qp();
  const double linf_primal = ScaledLInfNorm(x_ray, col_scale, sq.PrimalSharder());
  const double linf_dual = ScaledLInfNorm(y_ray, row_scale, sq.DualSharder());
  InfeasibilityInformation r;
  const Vec grad = PrimalGradientFromObjectiveProduct(sq, y_ray, ZeroVector(sq.PrimalSharder()), /*zero_objective=*/true);
  const ResidualNorms dr = DualResidualNorms(p, sq, col_scale, x_ray, grad, 0.0);
  const double dual_ray_objective = DualObjectiveBoundsTerm(sq, y_ray) + dr.objective_correction;
  if (linf_dual > 0) {
    r.dual_ray_objective = dual_ray_objective / linf_dual;
    r.max_dual_ray_infeasibility = dr.l_inf_residual / linf_dual;
  }
  const ResidualNorms pr = PrimalResidualNorms(sq, row_scale, x_ray, 0.0, /*homogeneous=*/true);
  const double sign_violation = sq.PrimalSharder().Max([&](int, int64_t b, int64_t e) {
    double mx = 0.0;
    for (int64_t j = b; j < e; ++j) {
      if (std::isfinite(qp.vlb[j])) mx = std::max(mx, -x_ray[j] * col_scale[j]);
      if (std::isfinite(qp.vub[j])) mx = std::max(mx, x_ray[j] * col_scale[j]);
    }
    return mx;
  });
  if (linf_primal > 0.0) {
    const Vec qx = ObjectiveProduct(sq, x_ray);
    r.primal_ray_quadratic_norm = LInfNorm(qx, sq.PrimalSharder()) / linf_primal;
    r.max_primal_ray_infeasibility = std::max(pr.l_inf_residual, sign_violation) / linf_primal;
    r.primal_ray_linear_objective = Dot(x_ray, qp.c, sq.PrimalSharder()) / linf_primal;
  }
  r.candidate_type = candidate_type;
  return r;
}

// iteration_stats.cc:533-565.
inline Vec ReducedCosts(const Params& p, const ShardedQp& sq, const Vec& x, const Vec& y, bool zero_objective = false) {
  Vec qx = zero_objective ? ZeroVector(sq.PrimalSharder()) : ObjectiveProduct(sq, x);
  Vec rc = PrimalGradientFromObjectiveProduct(sq, y, std::move(qx), zero_objective);
  const Qp& qp = sq.qp();
  sq.PrimalSharder().ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j)
      if (rc[j] != 0.0 && !HandlePrimalGradientTermAsReducedCost(p, rc[j], x[j], qp.vlb[j], qp.vub[j])) rc[j] = 0.0;
  });
  return rc;
}

inline const ConvergenceInformation* GetConvergenceInformation(const IterationStats& s, int type) {
  for (const auto& c : s.convergence_information) if (c.candidate_type == type) return &c;
  return nullptr;
}

// iteration_stats.cc:321-346,598-610.  parity unpinned (absl::Gaussian); see oracle_qp.h note.
inline double RandomProjection(const Vec& v, const Sharder& sh, std::mt19937& seed_gen) {
  std::vector<std::mt19937> seeds;
  for (int s = 0; s < sh.NumShards(); ++s) seeds.emplace_back(seed_gen());
  Vec dot(sh.NumShards(), 0.0), nrm(sh.NumShards(), 0.0);
  sh.ForEach([&](int s, int64_t b, int64_t e) {
    std::mt19937 rnd = seeds[s];
    std::normal_distribution<double> gauss(0.0, 1.0);
    double d = 0, n2 = 0;
    for (int64_t i = b; i < e; ++i) { const double g = gauss(rnd); d += g * v[i]; n2 += g * g; }
    dot[s] = d; nrm[s] = n2;
  });
  double d = 0, n2 = 0;
  for (int s = 0; s < sh.NumShards(); ++s) { d += dot[s]; n2 += nrm[s]; }
  return d / std::sqrt(n2);
}

// ---- termination.cc ----
struct BoundNorms { double l2_norm_primal_linear_objective = 0, l2_norm_constraint_bounds = 0, l_inf_norm_primal_linear_objective = 0, l_inf_norm_constraint_bounds = 0; };
struct TerminationReasonAndPointType { int reason; int type; };

inline DetailedCriteria EffectiveOptimalityCriteria(const TerminationCriteria& tc) {  // :126-159
  if (tc.has_detailed) return tc.detailed;
  double a = tc.eps_optimal_absolute, r = tc.eps_optimal_relative;
  if (tc.has_simple) { a = tc.simple_eps_abs; r = tc.simple_eps_rel; }
  return DetailedCriteria{a, r, a, r, a, r};
}

inline bool ObjectiveGapMet(const DetailedCriteria& oc, const ConvergenceInformation& s) {  // :26-41
  if (std::isinf(oc.eps_optimal_objective_gap_absolute) || std::isinf(oc.eps_optimal_objective_gap_relative)) return true;
  const double abs_obj = std::abs(s.primal_objective) + std::abs(s.dual_objective);
  const double gap = std::abs(s.primal_objective - s.dual_objective);
  return std::isfinite(abs_obj) && gap <= oc.eps_optimal_objective_gap_absolute + oc.eps_optimal_objective_gap_relative * abs_obj;
}

inline bool OptimalityCriteriaMet(const DetailedCriteria& oc, const ConvergenceInformation& s, int norm, const BoundNorms& bn) {  // :43-97
  double perr, pbase, derr, dbase;
  double pabs = oc.eps_optimal_primal_residual_absolute, dabs = oc.eps_optimal_dual_residual_absolute;
  switch (norm) {
    case NORM_L_INF: perr = s.l_inf_primal_residual; pbase = bn.l_inf_norm_constraint_bounds; derr = s.l_inf_dual_residual; dbase = bn.l_inf_norm_primal_linear_objective; break;
    case NORM_L2: perr = s.l2_primal_residual; pbase = bn.l2_norm_constraint_bounds; derr = s.l2_dual_residual; dbase = bn.l2_norm_primal_linear_objective; break;
    case NORM_L_INF_COMPONENTWISE: perr = s.l_inf_componentwise_primal_residual; pbase = 1.0; pabs = 0.0; derr = s.l_inf_componentwise_dual_residual; dbase = 1.0; dabs = 0.0; break;
    default: ORACLE_CHECK(false); return false;
  }
  const bool p_ok = std::isinf(oc.eps_optimal_primal_residual_absolute) || std::isinf(oc.eps_optimal_primal_residual_relative) ||
                    perr <= pabs + oc.eps_optimal_primal_residual_relative * pbase;
  const bool d_ok = std::isinf(oc.eps_optimal_dual_residual_absolute) || std::isinf(oc.eps_optimal_dual_residual_relative) ||
                    derr <= dabs + oc.eps_optimal_dual_residual_relative * dbase;
  return p_ok && d_ok && ObjectiveGapMet(oc, s);
}

inline bool PrimalInfeasibilityCriteriaMet(double eps, const InfeasibilityInformation& s) {  // :104-109
  if (s.dual_ray_objective <= 0.0) return false;
  return s.max_dual_ray_infeasibility / s.dual_ray_objective <= eps;
}
inline bool DualInfeasibilityCriteriaMet(double eps, const InfeasibilityInformation& s) {  // :113-122
  if (s.primal_ray_linear_objective >= 0.0) return false;
  return (s.max_primal_ray_infeasibility / -s.primal_ray_linear_objective <= eps) && (s.primal_ray_quadratic_norm / -s.primal_ray_linear_objective <= eps);
}

inline std::optional<TerminationReasonAndPointType> CheckSimpleTerminationCriteria(const TerminationCriteria& c, const IterationStats& s,
                                                                                  const volatile int* interrupt = nullptr) {  // :161-184
  if (s.iteration_number >= c.iteration_limit) return TerminationReasonAndPointType{TERMINATION_REASON_ITERATION_LIMIT, POINT_TYPE_NONE};
  if (s.cumulative_kkt_matrix_passes >= c.kkt_matrix_pass_limit) return TerminationReasonAndPointType{TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT, POINT_TYPE_NONE};
  if (s.cumulative_time_sec >= c.time_sec_limit) return TerminationReasonAndPointType{TERMINATION_REASON_TIME_LIMIT, POINT_TYPE_NONE};
  if (interrupt != nullptr && *interrupt != 0) return TerminationReasonAndPointType{TERMINATION_REASON_INTERRUPTED_BY_USER, POINT_TYPE_NONE};
  return std::nullopt;
}

inline std::optional<TerminationReasonAndPointType> CheckIterateTerminationCriteria(const TerminationCriteria& c, const IterationStats& s, const BoundNorms& bn,
                                                                                   bool force_numerical_termination = false) {  // :186-219
  const DetailedCriteria oc = EffectiveOptimalityCriteria(c);
  for (const auto& ci : s.convergence_information)
    if (OptimalityCriteriaMet(oc, ci, c.optimality_norm, bn)) return TerminationReasonAndPointType{TERMINATION_REASON_OPTIMAL, ci.candidate_type};
  for (const auto& ii : s.infeasibility_information) {
    if (PrimalInfeasibilityCriteriaMet(c.eps_primal_infeasible, ii)) return TerminationReasonAndPointType{TERMINATION_REASON_PRIMAL_INFEASIBLE, ii.candidate_type};
    if (DualInfeasibilityCriteriaMet(c.eps_dual_infeasible, ii)) return TerminationReasonAndPointType{TERMINATION_REASON_DUAL_INFEASIBLE, ii.candidate_type};
  }
  if (force_numerical_termination) return TerminationReasonAndPointType{TERMINATION_REASON_NUMERICAL_ERROR, POINT_TYPE_NONE};
  return std::nullopt;
}

inline BoundNorms BoundNormsFromProblemStats(const QpStats& s) {  // :221-228
  return {s.objective_vector_l2_norm, s.combined_bounds_l2_norm, s.objective_vector_abs_max, s.combined_bounds_max};
}
inline double EpsilonRatio(double a, double r) { return (a == r) ? 1.0 : a / r; }  // :230-237

struct RelativeConvergenceInformation { double relative_l_inf_primal_residual = 0, relative_l2_primal_residual = 0, relative_l_inf_dual_residual = 0, relative_l2_dual_residual = 0, relative_optimality_gap = 0; };
inline RelativeConvergenceInformation ComputeRelativeResiduals(const DetailedCriteria& oc, const ConvergenceInformation& s, const BoundNorms& bn) {  // :239-271
  const double rp = EpsilonRatio(oc.eps_optimal_primal_residual_absolute, oc.eps_optimal_primal_residual_relative);
  const double rd = EpsilonRatio(oc.eps_optimal_dual_residual_absolute, oc.eps_optimal_dual_residual_relative);
  const double rg = EpsilonRatio(oc.eps_optimal_objective_gap_absolute, oc.eps_optimal_objective_gap_relative);
  RelativeConvergenceInformation r;
  r.relative_l_inf_primal_residual = s.l_inf_primal_residual / (rp + bn.l_inf_norm_constraint_bounds);
  r.relative_l2_primal_residual = s.l2_primal_residual / (rp + bn.l2_norm_constraint_bounds);
  r.relative_l_inf_dual_residual = s.l_inf_dual_residual / (rd + bn.l_inf_norm_primal_linear_objective);
  r.relative_l2_dual_residual = s.l2_dual_residual / (rd + bn.l2_norm_primal_linear_objective);
  const double abs_obj = std::abs(s.primal_objective) + std::abs(s.dual_objective);
  r.relative_optimality_gap = (s.primal_objective - s.dual_objective) / (rg + abs_obj);
  return r;
}

}  // namespace oracle
#endif  // PDLP_ORACLE_STATS_H_
