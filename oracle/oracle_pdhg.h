// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// The PDHG driver: preprocessing (bound clean-up, stats, rescaling, initial step size
// and primal weight), the iteration loop with the three step rules, restarts,
// termination checks, feasibility polishing and result construction.
//   follows ortools/pdlp/primal_dual_hybrid_gradient.cc:71-96,717-1105,1204-1218,
//           1268-1286,1343-2774,2778-2831.
// Glop presolve (primal_dual_hybrid_gradient.cc:1107-1202) is host-only, needs
// ortools/glop and is not restated: use_glop=true is rejected as INVALID_PARAMETER.
#ifndef PDLP_ORACLE_PDHG_H_
#define PDLP_ORACLE_PDHG_H_

#include <chrono>

#include "oracle_tr.h"
#include "oracle_validate.h"

namespace oracle {

// primal_dual_hybrid_gradient.h:70-83 (same enumerator order).
enum class IterationType { kNormal = 0, kPrimalFeasibility = 1, kDualFeasibility = 2, kPresolveTermination = 3, kNormalTermination = 4, kFeasibilityPolishingTermination = 5 };

struct SolverResult { Vec primal_solution, dual_solution, reduced_costs; SolveLog solve_log; };
// Result of the bench-support step-loop timing (oracle_pdhg_impl.h:TimeSteps).
struct StepTiming { double seconds = 0, preprocessing_seconds = 0; int accepted_steps = 0, rejected_steps = 0; bool completed = false; };
struct IterationCallbackInfo { IterationType iteration_type; const TerminationCriteria* termination_criteria; const IterationStats* iteration_stats; const BoundNorms* bound_norms; };
using Callback = std::function<void(const IterationCallbackInfo&)>;

class WallTimer {
 public:
  void Start() { running_ = true; t0_ = Clock::now(); }
  void Stop() { if (running_) { acc_ += Elapsed(); running_ = false; } }
  double Get() const { return acc_ + (running_ ? Elapsed() : 0.0); }
 private:
  using Clock = std::chrono::steady_clock;
  double Elapsed() const { return std::chrono::duration<double>(Clock::now() - t0_).count(); }
  Clock::time_point t0_; double acc_ = 0.0; bool running_ = false;
};

inline int NumThreads(int num_threads, int num_shards, const Qp& qp) {  // :71-88
  int capped = num_threads;
  if (num_shards > 0) capped = std::min(capped, num_shards);
  const int64_t limit = std::max<int64_t>(qp.vlb.size(), qp.clb.size());
  capped = static_cast<int>(std::min<int64_t>(capped, limit));
  return std::max(capped, 1);
}
inline int NumShards(int num_threads, int num_shards) { return num_shards > 0 ? num_shards : (num_threads == 1 ? 1 : 4 * num_threads); }  // :93-96

inline SolverResult ErrorSolverResult(int reason, const std::string& msg) {  // :717-725
  SolverResult r;
  r.solve_log.termination_reason = reason;
  r.solve_log.termination_string = msg;
  return r;
}

inline SolverResult ConstructSolverResult(Vec x, Vec y, const IterationStats& stats, int reason, int output_type, SolveLog log) {  // :303-316
  log.iteration_count = stats.iteration_number;
  log.termination_reason = reason;
  log.solution_type = output_type;
  log.solve_time_sec = stats.cumulative_time_sec;
  log.solution_stats = stats;
  SolverResult r;
  r.primal_solution = std::move(x);
  r.dual_solution = std::move(y);
  r.solve_log = std::move(log);
  return r;
}

// :731-897 (warnings dropped; messages kept).
inline std::optional<SolverResult> CheckProblemStats(const QpStats& s, bool check_small) {
  const double kBig = 1e50, kSmall = 1e-50;
  using detail::SixDigits;
  auto err = [](const std::string& m) { return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, m); };
  if (std::isnan(s.constraint_matrix_l2_norm)) return err("Constraint matrix has a NAN.");
  if (s.constraint_matrix_abs_max > kBig) return err("Constraint matrix has a non-zero with absolute value " + SixDigits(s.constraint_matrix_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (s.constraint_matrix_col_min_l_inf_norm > 0 && s.constraint_matrix_col_min_l_inf_norm < kSmall)
    return err("Constraint matrix has a column with Linf norm " + SixDigits(s.constraint_matrix_col_min_l_inf_norm) + " which is less than limit of " + SixDigits(kSmall) + ".");
  if (s.constraint_matrix_row_min_l_inf_norm > 0 && s.constraint_matrix_row_min_l_inf_norm < kSmall)
    return err("Constraint matrix has a row with Linf norm " + SixDigits(s.constraint_matrix_row_min_l_inf_norm) + " which is less than limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.combined_bounds_l2_norm)) return err("Constraint bounds vector has a NAN.");
  if (s.combined_bounds_max > kBig) return err("Combined constraint bounds vector has a non-zero with absolute value " + SixDigits(s.combined_bounds_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_small && s.combined_bounds_min > 0 && s.combined_bounds_min < kSmall)
    return err("Combined constraint bounds vector has a non-zero with absolute value " + SixDigits(s.combined_bounds_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.combined_variable_bounds_l2_norm)) return err("Variable bounds vector has a NAN.");
  if (s.combined_variable_bounds_max > kBig) return err("Combined variable bounds vector has a non-zero with absolute value " + SixDigits(s.combined_variable_bounds_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_small && s.combined_variable_bounds_min > 0 && s.combined_variable_bounds_min < kSmall)
    return err("Combined variable bounds vector has a non-zero with absolute value " + SixDigits(s.combined_variable_bounds_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.objective_vector_l2_norm)) return err("Objective vector has a NAN.");
  if (s.objective_vector_abs_max > kBig) return err("Objective vector has a non-zero with absolute value " + SixDigits(s.objective_vector_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_small && s.objective_vector_abs_min > 0 && s.objective_vector_abs_min < kSmall)
    return err("Objective vector has a non-zero with absolute value " + SixDigits(s.objective_vector_abs_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.objective_matrix_l2_norm)) return err("Objective matrix has a NAN.");
  if (s.objective_matrix_abs_max > kBig) return err("Objective matrix has a non-zero with absolute value " + SixDigits(s.objective_matrix_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  return std::nullopt;
}

// :899-947.
inline std::optional<SolverResult> CheckInitialSolution(const ShardedQp& sq, const Vec& x0, const Vec& y0) {
  const double kBig = 1e50;
  using detail::SixDigits;
  auto err = [](const std::string& m) { return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION, m); };
  if ((int64_t)x0.size() != sq.PrimalSize()) return err("Initial primal solution has size " + std::to_string(x0.size()) + " which differs from problem primal size " + std::to_string(sq.PrimalSize()));
  if (std::isnan(Norm(x0, sq.PrimalSharder()))) return err("Initial primal solution has a NAN.");
  if (const double nrm = LInfNorm(x0, sq.PrimalSharder()); nrm > kBig) return err("Initial primal solution has an entry with absolute value " + SixDigits(nrm) + " which exceeds limit of " + SixDigits(kBig));
  if ((int64_t)y0.size() != sq.DualSize()) return err("Initial dual solution has size " + std::to_string(y0.size()) + " which differs from problem dual size " + std::to_string(sq.DualSize()));
  if (std::isnan(Norm(y0, sq.DualSharder()))) return err("Initial dual solution has a NAN.");
  if (const double nrm = LInfNorm(y0, sq.DualSharder()); nrm > kBig) return err("Initial dual solution has an entry with absolute value " + SixDigits(nrm) + " which exceeds limit of " + SixDigits(kBig));
  return std::nullopt;
}

// :1343-1412.
inline void SetActiveSetInformation(const ShardedQp& sq, const Vec& x, const Vec& y, const Vec& x_start, const Vec& y_start, PointMetadata& md) {
  const Qp& qp = sq.qp();
  auto p_active = [&](const Vec& v, int64_t j) { return v[j] > qp.vlb[j] && v[j] < qp.vub[j]; };
  auto d_active = [&](const Vec& v, int64_t i) { return v[i] != 0.0 || (qp.clb[i] == -kInf && qp.cub[i] == kInf); };
  md.has_active = true;
  md.active_primal_variable_count = (int64_t)sq.PrimalSharder().Sum([&](int, int64_t b, int64_t e) { int64_t c = 0; for (int64_t j = b; j < e; ++j) c += p_active(x, j); return (double)c; });
  md.active_primal_variable_change = (int64_t)sq.PrimalSharder().Sum([&](int, int64_t b, int64_t e) { int64_t c = 0; for (int64_t j = b; j < e; ++j) c += (p_active(x, j) != p_active(x_start, j)); return (double)c; });
  md.active_dual_variable_count = (int64_t)sq.DualSharder().Sum([&](int, int64_t b, int64_t e) { int64_t c = 0; for (int64_t i = b; i < e; ++i) c += d_active(y, i); return (double)c; });
  md.active_dual_variable_change = (int64_t)sq.DualSharder().Sum([&](int, int64_t b, int64_t e) { int64_t c = 0; for (int64_t i = b; i < e; ++i) c += (d_active(y, i) != d_active(y_start, i)); return (double)c; });
}

// The outer driver (primal_dual_hybrid_gradient.cc:318-502 decl, :711-1643 impl).
class PreprocessSolver {
 public:
  PreprocessSolver(Qp qp, const Params& params)
      : num_threads_(NumThreads(params.num_threads, params.num_shards, qp)),
        num_shards_(NumShards(num_threads_, params.num_shards)),
        sq_(std::move(qp), num_threads_, num_shards_) {}

  SolverResult PreprocessAndSolve(const Params& params, const Vec* x0, const Vec* y0, const volatile int* interrupt, Callback cb);
  std::optional<SolverResult> Prepare(const Params& params, const Vec* x0, const Vec* y0, SolveLog& log, Vec& x, Vec& y, double& step_size, double& primal_weight);
  bool TimeSteps(const Params& params, int warmup, int steps, StepTiming& out);

  // :1434-1518 (text logging dropped).
  std::optional<TerminationReasonAndPointType> UpdateIterationStatsAndCheckTermination(
      const Params& params, bool force_numerical_termination, const Vec& x_cur, const Vec& y_cur, const Vec* x_avg, const Vec* y_avg, const Vec* dx, const Vec* dy,
      const Vec& x_start, const Vec& y_start, const volatile int* interrupt, IterationType type, const IterationStats& full_stats, IterationStats& stats) {
    stats.convergence_information.emplace_back();
    stats.infeasibility_information.emplace_back();
    ComputeConvergenceAndInfeasibility(params, x_cur, y_cur, POINT_TYPE_CURRENT_ITERATE, &stats.convergence_information.back(), &stats.infeasibility_information.back());
    AddPointMetadata(params, x_cur, y_cur, POINT_TYPE_CURRENT_ITERATE, x_start, y_start, stats);
    if (x_avg != nullptr && y_avg != nullptr) {
      stats.convergence_information.emplace_back();
      stats.infeasibility_information.emplace_back();
      ComputeConvergenceAndInfeasibility(params, *x_avg, *y_avg, POINT_TYPE_AVERAGE_ITERATE, &stats.convergence_information.back(), &stats.infeasibility_information.back());
      AddPointMetadata(params, *x_avg, *y_avg, POINT_TYPE_AVERAGE_ITERATE, x_start, y_start, stats);
    }
    if (dx != nullptr && dy != nullptr) {
      stats.infeasibility_information.emplace_back();
      ComputeConvergenceAndInfeasibility(params, *dx, *dy, POINT_TYPE_ITERATE_DIFFERENCE, nullptr, &stats.infeasibility_information.back());
      AddPointMetadata(params, *dx, *dy, POINT_TYPE_ITERATE_DIFFERENCE, x_start, y_start, stats);
    }
    if (callback_) callback_({type, &params.termination_criteria, &stats, &original_bound_norms_});
    if (auto t = CheckIterateTerminationCriteria(params.termination_criteria, stats, original_bound_norms_, force_numerical_termination); t.has_value()) return t;
    return CheckSimpleTerminationCriteria(params.termination_criteria, full_stats, interrupt);
  }

  // :1520-1564 (no presolve branch).
  void ComputeConvergenceAndInfeasibility(const Params& params, const Vec& x, const Vec& y, int type, ConvergenceInformation* ci, InfeasibilityInformation* ii) const {
    const DetailedCriteria oc = EffectiveOptimalityCriteria(params.termination_criteria);
    const double pr = EpsilonRatio(oc.eps_optimal_primal_residual_absolute, oc.eps_optimal_primal_residual_relative);
    const double dr = EpsilonRatio(oc.eps_optimal_dual_residual_absolute, oc.eps_optimal_dual_residual_relative);
    if (ci) *ci = ComputeConvergenceInformation(params, sq_, col_scale_, row_scale_, x, y, pr, dr, type);
    if (ii) *ii = ComputeInfeasibilityInformation(params, sq_, col_scale_, row_scale_, x, y, type);
  }

  // :1568-1643.
  SolverResult ConstructOriginalSolverResult(const Params& params, SolverResult result) const {
    const bool zero_obj = result.solve_log.termination_reason == TERMINATION_REASON_PRIMAL_INFEASIBLE;
    result.reduced_costs = ReducedCosts(params, sq_, result.primal_solution, result.dual_solution, zero_obj);
    CwiseProductInPlace(col_scale_, sq_.PrimalSharder(), result.primal_solution);
    CwiseProductInPlace(row_scale_, sq_.DualSharder(), result.dual_solution);
    CwiseQuotientInPlace(col_scale_, sq_.PrimalSharder(), result.reduced_costs);
    IterationType type = IterationType::kNormalTermination;
    if (result.solve_log.solution_type == POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION) type = IterationType::kFeasibilityPolishingTermination;
    else if (result.solve_log.solution_type == POINT_TYPE_PRESOLVER_SOLUTION) type = IterationType::kPresolveTermination;
    if (callback_) callback_({type, &params.termination_criteria, &result.solve_log.solution_stats, &original_bound_norms_});
    return result;
  }

  ShardedQp& sq() { return sq_; }
  const ShardedQp& sq() const { return sq_; }
  const BoundNorms& OriginalBoundNorms() const { return original_bound_norms_; }
  const Vec& col_scale() const { return col_scale_; }
  const Vec& row_scale() const { return row_scale_; }

 private:
  // :1414-1432.
  void AddPointMetadata(const Params& params, const Vec& x, const Vec& y, int type, const Vec& x_start, const Vec& y_start, IterationStats& stats) const {
    PointMetadata md;
    md.point_type = type;
    for (int seed : params.random_projection_seeds) {
      std::mt19937 gen(seed);
      md.random_primal_projections.push_back(RandomProjection(x, sq_.PrimalSharder(), gen));
      md.random_dual_projections.push_back(RandomProjection(y, sq_.DualSharder(), gen));
    }
    if (type != POINT_TYPE_ITERATE_DIFFERENCE) SetActiveSetInformation(sq_, x, y, x_start, y_start, md);
    stats.point_metadata.push_back(std::move(md));
  }

  const int num_threads_, num_shards_;
  BoundNorms original_bound_norms_;
  ShardedQp sq_;
  Vec col_scale_, row_scale_;
  Callback callback_;
};

enum class InnerStepOutcome { kSuccessful, kForceNumericalTermination };

// The PDHG loop (primal_dual_hybrid_gradient.cc:504-709 decl, :1645-2774 impl).
class Solver {
 public:
  Solver(const Params& params, Vec x0, Vec y0, double step_size, double primal_weight, PreprocessSolver* pre)
      : params_(params), x_(std::move(x0)), y_(std::move(y0)), x_avg_(&pre->sq().PrimalSharder()), y_avg_(&pre->sq().DualSharder()),
        step_size_(step_size), primal_weight_(primal_weight), pre_(pre) {}

  SolverResult Solve(IterationType type, const volatile int* interrupt, SolveLog log);
  void TimeSteps(int warmup, int steps, StepTiming& out);

 private:
  struct Next { Vec value, delta; };
  static constexpr double kDivergentMovement = 1.0e100;
  const ShardedQp& sq() const { return pre_->sq(); }
  const Qp& qp() const { return pre_->sq().qp(); }

  Next ComputeNextPrimalSolution(double tau) const {  // :1659-1705
    const int64_t n = sq().PrimalSize();
    Next r{Vec(n), Vec(n)};
    const Qp& q = qp();
    sq().PrimalSharder().ForEach([&](int, int64_t b, int64_t e) {
      if (!q.IsLp()) {
        for (int64_t j = b; j < e; ++j) {
          const double scale = tau * (*q.q)[j] + 1.0;
          const double t = (x_[j] - tau * (q.c[j] - kty_[j])) / scale;
          r.value[j] = std::max(std::min(t, q.vub[j]), q.vlb[j]);
        }
      } else {
        for (int64_t j = b; j < e; ++j) {
          const double t = x_[j] - tau * (q.c[j] - kty_[j]);
          r.value[j] = std::max(std::min(t, q.vub[j]), q.vlb[j]);
        }
      }
      for (int64_t j = b; j < e; ++j) r.delta[j] = r.value[j] - x_[j];
    });
    return r;
  }

  Next ComputeNextDualSolution(double sigma, double extrapolation, const Next& np) const {  // :1707-1748
    const int64_t m = sq().DualSize();
    Next r{Vec(m), Vec(m)};
    const Qp& q = qp();
    Vec xbar(sq().PrimalSize());
    sq().PrimalSharder().ForEach([&](int, int64_t b, int64_t e) { for (int64_t j = b; j < e; ++j) xbar[j] = np.value[j] + extrapolation * np.delta[j]; });
    const Csc& kt = sq().Kt();
    sq().KtSharder().ForEach([&](int, int64_t b, int64_t e) {
      for (int64_t i = b; i < e; ++i) {
        double kx = 0.0;
        for (int64_t k = kt.ptr[i]; k < kt.ptr[i + 1]; ++k) kx += kt.val[k] * xbar[kt.idx[k]];
        const double temp = y_[i] - sigma * kx;
        r.value[i] = std::max(std::min(0.0, temp + sigma * q.cub[i]), temp + sigma * q.clb[i]);
        r.delta[i] = r.value[i] - y_[i];
      }
    });
    return r;
  }

  double ComputeMovement(const Vec& dx, const Vec& dy) const {  // :1750-1759
    return (0.5 * primal_weight_) * SquaredNorm(dx, sq().PrimalSharder()) + (0.5 / primal_weight_) * SquaredNorm(dy, sq().DualSharder());
  }
  double ComputeNonlinearity(const Vec& dx, const Vec& next_kty) const {  // :1761-1771
    return sq().PrimalSharder().Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t j = b; j < e; ++j) s += dx[j] * (next_kty[j] - kty_[j]); return -s; });
  }

  IterationStats CreateSimpleIterationStats(int restart_used) const {  // :1773-1793
    IterationStats s;
    const double per_rejected = params_.linesearch_rule == MALITSKY_POCK_LINESEARCH_RULE ? 0.5 : 1.0;
    s.iteration_number = iterations_completed_;
    s.cumulative_rejected_steps = num_rejected_steps_;
    s.cumulative_kkt_matrix_passes = iterations_completed_ + per_rejected * num_rejected_steps_;
    s.cumulative_time_sec = preprocessing_time_sec_ + timer_.Get();
    s.restart_used = restart_used;
    s.step_size = step_size_;
    s.primal_weight = primal_weight_;
    return s;
  }

  double DistanceTraveledFromLastStart(const Vec& x, const Vec& y) const {  // :1795-1804
    return std::sqrt((0.5 * primal_weight_) * SquaredDistance(x, x_start_, sq().PrimalSharder()) + (0.5 / primal_weight_) * SquaredDistance(y, y_start_, sq().DualSharder()));
  }
  LocalizedLagrangianBounds BoundsAtCurrent() const {  // :1806-1816
    return ComputeLocalizedLagrangianBounds(sq(), x_, y_, PrimalDualNorm::kEuclideanNorm, primal_weight_, DistanceTraveledFromLastStart(x_, y_), nullptr, &kty_,
                                            params_.use_diagonal_qp_trust_region_solver, params_.diagonal_qp_trust_region_solver_tolerance);
  }
  LocalizedLagrangianBounds BoundsAtAverage() const {  // :1818-1834
    const Vec xa = PrimalAverage(), ya = DualAverage();
    return ComputeLocalizedLagrangianBounds(sq(), xa, ya, PrimalDualNorm::kEuclideanNorm, primal_weight_, DistanceTraveledFromLastStart(xa, ya), nullptr, nullptr,
                                            params_.use_diagonal_qp_trust_region_solver, params_.diagonal_qp_trust_region_solver_tolerance);
  }
  static bool AverageHasBetterPotential(const LocalizedLagrangianBounds& a, const LocalizedLagrangianBounds& c) {  // :1836-1843
    return BoundGap(a) / (a.radius * a.radius) < BoundGap(c) / (c.radius * c.radius);
  }
  static double NormalizedGap(const LocalizedLagrangianBounds& b) { return BoundGap(b) / b.radius; }  // :1845-1850
  bool ShouldDoAdaptiveRestartHeuristic(double candidate_gap) const {  // :1853-1867
    const double ratio = candidate_gap / normalized_gap_at_last_restart_;
    if (ratio < params_.sufficient_reduction_for_restart) return true;
    if (ratio < params_.necessary_reduction_for_restart && candidate_gap > normalized_gap_at_last_trial_) return true;
    return false;
  }
  int DetermineDistanceBasedRestartChoice() const {  // :1869-1902
    if (x_avg_.NumTerms() == 0) return RESTART_CHOICE_NO_RESTART;
    if (distance_info_.length_of_last_restart_period == 0) return RESTART_CHOICE_RESTART_TO_AVERAGE;
    const int len = x_avg_.NumTerms();
    const double dist_avg = DistanceTraveledFromLastStart(x_avg_.ComputeAverage(), y_avg_.ComputeAverage());
    if ((dist_avg / len) < params_.sufficient_reduction_for_restart * (distance_info_.distance_moved_last_restart_period / distance_info_.length_of_last_restart_period)) {
      return AverageHasBetterPotential(BoundsAtAverage(), BoundsAtCurrent()) ? RESTART_CHOICE_RESTART_TO_AVERAGE : RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
    }
    return RESTART_CHOICE_NO_RESTART;
  }
  int ChooseRestartToApply(bool is_major);  // :1904-1965
  Vec PrimalAverage() const { return x_avg_.HasNonzeroWeight() ? x_avg_.ComputeAverage() : x_; }  // :1967-1973
  Vec DualAverage() const { return y_avg_.HasNonzeroWeight() ? y_avg_.ComputeAverage() : y_; }    // :1975-1981
  double ComputeNewPrimalWeight() const {  // :1983-2010
    const double dp = Distance(x_, x_start_, sq().PrimalSharder());
    const double dd = Distance(y_, y_start_, sq().DualSharder());
    constexpr double kTol = 1.0e-10;
    if (dp <= kTol || dp >= 1.0 / kTol || dd <= kTol || dd >= 1.0 / kTol) return primal_weight_;
    const double th = params_.primal_weight_update_smoothing;
    return std::exp(th * std::log(dd / dp) + (1.0 - th) * std::log(primal_weight_));
  }
  SolverResult PickSolutionAndConstructSolverResult(Vec x, Vec y, const IterationStats& stats, int reason, int type, SolveLog log) const {  // :2012-2040
    switch (type) {
      case POINT_TYPE_CURRENT_ITERATE: x = x_; y = y_; break;
      case POINT_TYPE_ITERATE_DIFFERENCE: x = dx_; y = dy_; break;
      case POINT_TYPE_AVERAGE_ITERATE: case POINT_TYPE_PRESOLVER_SOLUTION: break;
      default: type = POINT_TYPE_AVERAGE_ITERATE; break;
    }
    return ConstructSolverResult(std::move(x), std::move(y), stats, reason, type, std::move(log));
  }
  void ApplyRestartChoice(int restart);  // :2042-2094
  std::optional<SolverResult> MajorIterationAndTerminationCheck(IterationType type, bool force_numerical, const volatile int* interrupt, const IterationStats& polishing_work, SolveLog& log);
  void ResetAverageToCurrent() { x_avg_.Clear(); y_avg_.Clear(); x_avg_.Add(x_, 1.0); y_avg_.Add(y_, 1.0); }  // :2192-2197
  InnerStepOutcome TakeMalitskyPockStep();
  InnerStepOutcome TakeAdaptiveStep();
  InnerStepOutcome TakeConstantSizeStep();
  IterationStats TotalWorkSoFar(const SolveLog& log) const;
  std::optional<SolverResult> TryFeasibilityPolishing(int iteration_limit, const volatile int* interrupt, SolveLog& log);
  SolverResult TryPrimalPolishing(Vec x_start, int iteration_limit, const volatile int* interrupt, SolveLog& log);
  SolverResult TryDualPolishing(Vec y_start, int iteration_limit, const volatile int* interrupt, SolveLog& log);

  const Params params_;
  Vec x_, y_, dx_, dy_;
  WeightedAverage x_avg_, y_avg_;
  double step_size_, primal_weight_;
  PreprocessSolver* pre_;
  double ratio_last_two_step_sizes_ = 1.0;
  double normalized_gap_at_last_trial_ = kInf, normalized_gap_at_last_restart_ = kInf;
  double preprocessing_time_sec_ = 0.0;
  WallTimer timer_;
  int iterations_completed_ = 0, num_rejected_steps_ = 0;
  Vec kty_;  // cache of K^T y
  Vec x_start_, y_start_;
  struct { double distance_moved_last_restart_period = kInf; int length_of_last_restart_period = 1; } distance_info_;
};

}  // namespace oracle
#include "oracle_pdhg_impl.h"
#endif  // PDLP_ORACLE_PDHG_H_
