// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
// Out-of-line members of oracle::PreprocessSolver / oracle::Solver.
//   follows ortools/pdlp/primal_dual_hybrid_gradient.cc:949-1105,1904-1965,2042-2774.
#ifndef PDLP_ORACLE_PDHG_IMPL_H_
#define PDLP_ORACLE_PDHG_IMPL_H_

namespace oracle {

// :949-1105 (presolve branch :988-1033 not restated).  Prepare() is everything before the Solver is
// built: bound clean-up, statistics, input checks, projection of the start, rescaling, initial step
// size and primal weight; PreprocessAndSolve() then runs the loop.
inline std::optional<SolverResult> PreprocessSolver::Prepare(const Params& params, const Vec* x0, const Vec* y0, SolveLog& log, Vec& x, Vec& y, double& step_size,
                                                             double& primal_weight) {
  if (sq_.qp().name.has_value()) log.instance_name = *sq_.qp().name;
  sq_.ReplaceLargeConstraintBoundsWithInfinity(params.infinite_constraint_bound_threshold);
  log.original_problem_stats = ComputeStats(sq_);
  log.has_original_stats = true;
  if (auto r = CheckProblemStats(log.original_problem_stats, params.use_glop); r.has_value()) return r;
  if (x0 != nullptr && y0 != nullptr) {
    if (auto r = CheckInitialSolution(sq_, *x0, *y0); r.has_value()) return r;
  }
  original_bound_norms_ = BoundNormsFromProblemStats(log.original_problem_stats);

  if (x0 != nullptr && y0 != nullptr) { x = *x0; y = *y0; }
  else { SetZero(sq_.PrimalSharder(), x); SetZero(sq_.DualSharder(), y); }
  ProjectToPrimalVariableBounds(sq_, x);
  ProjectToDualVariableBounds(sq_, y);

  // ComputeAndApplyRescaling :1204-1218.
  ScalingVectors sc = ApplyRescaling(params.l_inf_ruiz_iterations, params.l2_norm_rescaling, sq_);
  row_scale_ = std::move(sc.row);
  col_scale_ = std::move(sc.col);
  CwiseQuotientInPlace(col_scale_, sq_.PrimalSharder(), x);
  CwiseQuotientInPlace(row_scale_, sq_.DualSharder(), y);
  log.preprocessed_problem_stats = ComputeStats(sq_);
  log.has_preprocessed_stats = true;

  step_size = 0.0;
  if (params.linesearch_rule == CONSTANT_STEP_SIZE_RULE) {  // :1060-1076
    std::mt19937 random(1);
    const auto est = EstimateMaximumSingularValue(sq_, nullptr, nullptr, 0.2, 0.0005, random);
    const double inv = est.singular_value / (1.0 - est.estimated_relative_error);
    step_size = inv > 0.0 ? 1.0 / inv : 1.0;
  } else {
    step_size = 1.0 / std::max(1.0e-20, log.preprocessed_problem_stats.constraint_matrix_abs_max);
  }
  step_size *= params.initial_step_size_scaling;

  // InitialPrimalWeight :1268-1286.
  primal_weight = 1.0;
  const double c_norm = log.preprocessed_problem_stats.objective_vector_l2_norm, b_norm = log.preprocessed_problem_stats.combined_bounds_l2_norm;
  if (params.has_initial_primal_weight) primal_weight = params.initial_primal_weight;
  else if (c_norm > 0.0 && b_norm > 0.0) primal_weight = c_norm / b_norm;
  return std::nullopt;
}

inline SolverResult PreprocessSolver::PreprocessAndSolve(const Params& params, const Vec* x0, const Vec* y0, const volatile int* interrupt, Callback cb) {
  WallTimer timer;
  timer.Start();
  SolveLog log;
  callback_ = std::move(cb);
  Vec x, y;
  double step_size = 0.0, primal_weight = 1.0;
  if (auto r = Prepare(params, x0, y0, log, x, y, step_size, primal_weight); r.has_value()) return *r;
  Solver solver(params, x, y, step_size, primal_weight, this);
  log.preprocessing_time_sec = timer.Get();
  SolverResult result = solver.Solve(IterationType::kNormal, interrupt, std::move(log));
  return ConstructOriginalSolverResult(params, std::move(result));
}

// Bench support (bench.py cpu_baseline / --impl reference): the steady-state cost of the step loop
// alone - `warmup` untimed + `steps` timed iterations of the configured step rule
// (TakeAdaptiveStep / TakeMalitskyPockStep / TakeConstantSizeStep, rejected trials included) after
// the reference's preprocessing, with no termination checks and no restarts: the same unit as the
// device's `pdlp_b200_trial_steps`.  Returns false when preprocessing rejects the problem.
inline bool PreprocessSolver::TimeSteps(const Params& params, int warmup, int steps, StepTiming& out) {
  WallTimer timer;
  timer.Start();
  SolveLog log;
  Vec x, y;
  double step_size = 0.0, primal_weight = 1.0;
  if (Prepare(params, nullptr, nullptr, log, x, y, step_size, primal_weight).has_value()) return false;
  Solver solver(params, x, y, step_size, primal_weight, this);
  out.preprocessing_seconds = timer.Get();
  solver.TimeSteps(warmup, steps, out);
  return true;
}

inline void Solver::TimeSteps(int warmup, int steps, StepTiming& out) {
  x_start_ = CloneVector(x_, sq().PrimalSharder());
  y_start_ = CloneVector(y_, sq().DualSharder());
  ratio_last_two_step_sizes_ = 1;
  kty_ = TransposedMatVec(qp().K, y_, sq().KSharder());
  num_rejected_steps_ = 0;
  auto run = [&](int count) {
    for (int i = 0; i < count; ++i, ++iterations_completed_) {
      InnerStepOutcome outcome;
      switch (params_.linesearch_rule) {
        case MALITSKY_POCK_LINESEARCH_RULE: outcome = TakeMalitskyPockStep(); break;
        case CONSTANT_STEP_SIZE_RULE: outcome = TakeConstantSizeStep(); break;
        default: outcome = TakeAdaptiveStep(); break;
      }
      if (outcome == InnerStepOutcome::kForceNumericalTermination) return false;
    }
    return true;
  };
  iterations_completed_ = 0;
  out.completed = run(warmup);
  const int rejected_before = num_rejected_steps_;
  WallTimer timer;
  timer.Start();
  out.completed = out.completed && run(steps);
  out.seconds = timer.Get();
  out.accepted_steps = iterations_completed_ - warmup;
  out.rejected_steps = num_rejected_steps_ - rejected_before;
}

inline int Solver::ChooseRestartToApply(bool is_major) {  // :1904-1965
  if (!x_avg_.HasNonzeroWeight() && !y_avg_.HasNonzeroWeight()) return RESTART_CHOICE_NO_RESTART;
  const int restart_length = x_avg_.NumTerms();
  if (restart_length >= iterations_completed_ / 2 && params_.restart_strategy == ADAPTIVE_HEURISTIC) {
    return AverageHasBetterPotential(BoundsAtAverage(), BoundsAtCurrent()) ? RESTART_CHOICE_RESTART_TO_AVERAGE : RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
  }
  if (!is_major) return RESTART_CHOICE_NO_RESTART;
  switch (params_.restart_strategy) {
    case NO_RESTARTS: return RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
    case EVERY_MAJOR_ITERATION: return RESTART_CHOICE_RESTART_TO_AVERAGE;
    case ADAPTIVE_HEURISTIC: {
      const LocalizedLagrangianBounds at_avg = BoundsAtAverage();
      const LocalizedLagrangianBounds at_cur = BoundsAtCurrent();
      double gap; int choice;
      if (AverageHasBetterPotential(at_avg, at_cur)) { gap = NormalizedGap(at_avg); choice = RESTART_CHOICE_RESTART_TO_AVERAGE; }
      else { gap = NormalizedGap(at_cur); choice = RESTART_CHOICE_WEIGHTED_AVERAGE_RESET; }
      if (ShouldDoAdaptiveRestartHeuristic(gap)) return choice;
      normalized_gap_at_last_trial_ = gap;
      return RESTART_CHOICE_NO_RESTART;
    }
    case ADAPTIVE_DISTANCE_BASED: return DetermineDistanceBasedRestartChoice();
    default: ORACLE_CHECK(false); return RESTART_CHOICE_UNSPECIFIED;
  }
}

inline void Solver::ApplyRestartChoice(int restart) {  // :2042-2094
  switch (restart) {
    case RESTART_CHOICE_UNSPECIFIED: case RESTART_CHOICE_NO_RESTART: return;
    case RESTART_CHOICE_WEIGHTED_AVERAGE_RESET: break;
    case RESTART_CHOICE_RESTART_TO_AVERAGE:
      x_ = x_avg_.ComputeAverage();
      y_ = y_avg_.ComputeAverage();
      kty_ = TransposedMatVec(qp().K, y_, sq().KSharder());
      break;
  }
  primal_weight_ = ComputeNewPrimalWeight();
  ratio_last_two_step_sizes_ = 1;
  if (params_.restart_strategy == ADAPTIVE_HEURISTIC) {
    const LocalizedLagrangianBounds b = BoundsAtCurrent();
    normalized_gap_at_last_restart_ = BoundGap(b) / b.radius;
    normalized_gap_at_last_trial_ = kInf;
  } else if (params_.restart_strategy == ADAPTIVE_DISTANCE_BASED) {
    distance_info_.distance_moved_last_restart_period = DistanceTraveledFromLastStart(x_, y_);
    distance_info_.length_of_last_restart_period = x_avg_.NumTerms();
  }
  x_avg_.Clear();
  y_avg_.Clear();
  AssignVector(x_, sq().PrimalSharder(), x_start_);
  AssignVector(y_, sq().DualSharder(), y_start_);
}

inline IterationStats AddWorkStats(IterationStats s, const IterationStats& extra) {  // :2100-2113
  s.iteration_number += extra.iteration_number;
  s.cumulative_kkt_matrix_passes += extra.cumulative_kkt_matrix_passes;
  s.cumulative_rejected_steps += extra.cumulative_rejected_steps;
  s.cumulative_time_sec += extra.cumulative_time_sec;
  return s;
}
inline IterationStats WorkFromFeasibilityPolishing(const SolveLog& log) {  // :2118-2126
  IterationStats r;
  for (const auto& d : log.feasibility_polishing_details) r = AddWorkStats(std::move(r), d.solution_stats);
  return r;
}

inline std::optional<SolverResult> Solver::MajorIterationAndTerminationCheck(IterationType type, bool force_numerical, const volatile int* interrupt,
                                                                            const IterationStats& polishing_work, SolveLog& log) {  // :2128-2190
  const int cycle = iterations_completed_ % params_.major_iteration_frequency;
  const bool is_major = cycle == 0 && iterations_completed_ > 0;
  const int restart = force_numerical ? RESTART_CHOICE_NO_RESTART : ChooseRestartToApply(is_major);
  IterationStats stats = CreateSimpleIterationStats(restart);
  const IterationStats full_work = AddWorkStats(stats, polishing_work);
  const bool check = cycle % params_.termination_check_frequency == 0 ||
                     CheckSimpleTerminationCriteria(params_.termination_criteria, full_work, interrupt).has_value() || force_numerical;
  if (check) {
    Vec xa = PrimalAverage(), ya = DualAverage();
    const auto reason = pre_->UpdateIterationStatsAndCheckTermination(
        params_, force_numerical, x_, y_, x_avg_.HasNonzeroWeight() ? &xa : nullptr, y_avg_.HasNonzeroWeight() ? &ya : nullptr,
        dx_.size() > 0 ? &dx_ : nullptr, dy_.size() > 0 ? &dy_ : nullptr, x_start_, y_start_, interrupt, type, full_work, stats);
    if (params_.record_iteration_stats) log.iteration_stats.push_back(stats);
    if (reason.has_value()) {
      const IterationStats final_stats = AddWorkStats(stats, polishing_work);
      return PickSolutionAndConstructSolverResult(std::move(xa), std::move(ya), final_stats, reason->reason, reason->type, std::move(log));
    }
  } else if (params_.record_iteration_stats) {
    log.iteration_stats.push_back(stats);
  }
  ApplyRestartChoice(restart);
  return std::nullopt;
}

inline InnerStepOutcome Solver::TakeMalitskyPockStep() {  // :2211-2293
  InnerStepOutcome outcome = InnerStepOutcome::kSuccessful;
  const double tau = step_size_ / primal_weight_;
  Next np = ComputeNextPrimalSolution(tau);
  const double dilating = 1 + (params_.step_size_interpolation * (std::sqrt(1 + ratio_last_two_step_sizes_) - 1));
  double new_tau = tau * dilating;
  const double dual_weight = primal_weight_ * primal_weight_;
  int inner = 0;
  for (bool accepted = false; !accepted; ++inner) {
    if (inner >= 60) { ResetAverageToCurrent(); outcome = InnerStepOutcome::kForceNumericalTermination; break; }
    const double new_ratio = new_tau / tau;
    Next nd = ComputeNextDualSolution(dual_weight * new_tau, new_ratio, np);
    Vec next_kty = TransposedMatVec(qp().K, nd.value, sq().KSharder());
    const double dy_norm = Norm(nd.delta, sq().DualSharder());
    const double dkty_norm = Distance(kty_, next_kty, sq().PrimalSharder());
    if (primal_weight_ * new_tau * dkty_norm <= params_.linesearch_contraction_factor * dy_norm) {
      step_size_ = new_tau * primal_weight_;
      ratio_last_two_step_sizes_ = new_ratio;
      if (!x_avg_.HasNonzeroWeight()) x_avg_.Add(x_, new_tau * new_ratio);
      x_ = std::move(np.value);
      y_ = std::move(nd.value);
      kty_ = std::move(next_kty);
      x_avg_.Add(x_, new_tau);
      y_avg_.Add(y_, new_tau);
      const double movement = ComputeMovement(np.delta, nd.delta);
      if (movement == 0.0) { ResetAverageToCurrent(); outcome = InnerStepOutcome::kForceNumericalTermination; }
      else if (movement > kDivergentMovement) outcome = InnerStepOutcome::kForceNumericalTermination;
      dx_ = std::move(np.delta);
      dy_ = std::move(nd.delta);
      break;
    }
    new_tau = params_.step_size_downscaling_factor * new_tau;
  }
  num_rejected_steps_ += inner;
  return outcome;
}

inline InnerStepOutcome Solver::TakeAdaptiveStep() {  // :2295-2376
  InnerStepOutcome outcome = InnerStepOutcome::kSuccessful;
  int inner = 0;
  for (bool accepted = false; !accepted; ++inner) {
    if (inner >= 60) { ResetAverageToCurrent(); outcome = InnerStepOutcome::kForceNumericalTermination; break; }
    const double tau = step_size_ / primal_weight_;
    const double sigma = step_size_ * primal_weight_;
    Next np = ComputeNextPrimalSolution(tau);
    Next nd = ComputeNextDualSolution(sigma, 1.0, np);
    const double movement = ComputeMovement(np.delta, nd.delta);
    if (movement == 0.0) { ResetAverageToCurrent(); outcome = InnerStepOutcome::kForceNumericalTermination; break; }
    if (movement > kDivergentMovement) { outcome = InnerStepOutcome::kForceNumericalTermination; break; }
    Vec next_kty = TransposedMatVec(qp().K, nd.value, sq().KSharder());
    const double nonlinearity = ComputeNonlinearity(np.delta, next_kty);
    const double limit = nonlinearity > 0 ? movement / nonlinearity : kInf;
    if (step_size_ <= limit) {
      x_ = std::move(np.value);
      y_ = std::move(nd.value);
      kty_ = std::move(next_kty);
      dx_ = std::move(np.delta);
      dy_ = std::move(nd.delta);
      x_avg_.Add(x_, step_size_);
      y_avg_.Add(y_, step_size_);
      accepted = true;
    }
    const double total_steps = num_rejected_steps_ + inner + iterations_completed_ + 1;
    const double first = std::isinf(limit) ? limit : (1 - std::pow(total_steps + 1.0, -params_.step_size_reduction_exponent)) * limit;
    const double second = (1 + std::pow(total_steps + 1.0, -params_.step_size_growth_exponent)) * step_size_;
    step_size_ = std::min(first, second);
  }
  num_rejected_steps_ += inner - 1;
  return outcome;
}

inline InnerStepOutcome Solver::TakeConstantSizeStep() {  // :2378-2406
  const double tau = step_size_ / primal_weight_, sigma = step_size_ * primal_weight_;
  Next np = ComputeNextPrimalSolution(tau);
  Next nd = ComputeNextDualSolution(sigma, 1.0, np);
  const double movement = ComputeMovement(np.delta, nd.delta);
  if (movement == 0.0) { ResetAverageToCurrent(); return InnerStepOutcome::kForceNumericalTermination; }
  if (movement > kDivergentMovement) return InnerStepOutcome::kForceNumericalTermination;
  Vec next_kty = TransposedMatVec(qp().K, nd.value, sq().KSharder());
  x_ = std::move(np.value);
  y_ = std::move(nd.value);
  kty_ = std::move(next_kty);
  dx_ = std::move(np.delta);
  dy_ = std::move(nd.delta);
  x_avg_.Add(x_, step_size_);
  y_avg_.Add(y_, step_size_);
  return InnerStepOutcome::kSuccessful;
}

inline IterationStats Solver::TotalWorkSoFar(const SolveLog& log) const {  // :2408-2413
  return AddWorkStats(CreateSimpleIterationStats(RESTART_CHOICE_NO_RESTART), WorkFromFeasibilityPolishing(log));
}

inline FeasibilityPolishingDetails BuildFeasibilityPolishingDetails(int phase, int main_iterations, const Params& params, const SolveLog& log) {  // :2415-2431
  FeasibilityPolishingDetails d;
  d.polishing_phase_type = phase;
  d.main_iteration_count = main_iterations;
  d.detailed_criteria = params.termination_criteria.detailed;  // the part of `params` the phase changes
  d.iteration_limit = params.termination_criteria.iteration_limit;
  d.termination_reason = log.termination_reason;
  d.iteration_count = log.iteration_count;
  d.solve_time_sec = log.solve_time_sec;
  d.solution_stats = log.solution_stats;
  d.solution_type = log.solution_type;
  d.iteration_stats = log.iteration_stats;
  return d;
}
inline bool TerminationReasonIsWorkLimit(int r) {  // :2435-2440
  return r == TERMINATION_REASON_ITERATION_LIMIT || r == TERMINATION_REASON_TIME_LIMIT || r == TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT || r == TERMINATION_REASON_INTERRUPTED_BY_USER;
}

inline std::optional<SolverResult> Solver::TryFeasibilityPolishing(int iteration_limit, const volatile int* interrupt, SolveLog& log) {  // :2442-2552
  const DetailedCriteria oc = EffectiveOptimalityCriteria(params_.termination_criteria);
  Vec xa = PrimalAverage(), ya = DualAverage();
  ConvergenceInformation first;
  pre_->ComputeConvergenceAndInfeasibility(params_, xa, ya, POINT_TYPE_AVERAGE_ITERATE, &first, nullptr);
  if (!ObjectiveGapMet(oc, first)) return std::nullopt;
  SolverResult pr = TryPrimalPolishing(std::move(xa), iteration_limit, interrupt, log);
  if (TerminationReasonIsWorkLimit(pr.solve_log.termination_reason)) return std::nullopt;
  if (pr.solve_log.termination_reason != TERMINATION_REASON_OPTIMAL) return std::nullopt;
  SolverResult dr = TryDualPolishing(std::move(ya), iteration_limit, interrupt, log);
  if (TerminationReasonIsWorkLimit(dr.solve_log.termination_reason)) return std::nullopt;
  if (dr.solve_log.termination_reason != TERMINATION_REASON_OPTIMAL) return std::nullopt;
  IterationStats full = TotalWorkSoFar(log);
  full.convergence_information.emplace_back();
  pre_->ComputeConvergenceAndInfeasibility(params_, pr.primal_solution, dr.dual_solution, POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION, &full.convergence_information.back(), nullptr);
  const auto earned = CheckIterateTerminationCriteria(params_.termination_criteria, full, pre_->OriginalBoundNorms(), false);
  if (earned.has_value()) {
    return ConstructSolverResult(std::move(pr.primal_solution), std::move(dr.dual_solution), full, earned->reason, POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION, log);
  }
  return std::nullopt;
}

inline TerminationCriteria ReduceWorkLimitsByPreviousWork(TerminationCriteria c, int iteration_limit, const IterationStats& prev) {  // :2554-2566
  c.iteration_limit = std::max(0, std::min(iteration_limit, c.iteration_limit - prev.iteration_number));
  c.kkt_matrix_pass_limit = std::max(0.0, c.kkt_matrix_pass_limit - prev.cumulative_kkt_matrix_passes);
  c.time_sec_limit = std::max(0.0, c.time_sec_limit - prev.cumulative_time_sec);
  return c;
}

inline SolverResult Solver::TryPrimalPolishing(Vec x_start, int iteration_limit, const volatile int* interrupt, SolveLog& log) {  // :2568-2614
  Params p = params_;
  p.termination_criteria = ReduceWorkLimitsByPreviousWork(params_.termination_criteria, iteration_limit, TotalWorkSoFar(log));
  Vec objective;
  SetZero(sq().PrimalSharder(), objective);
  pre_->sq().SwapObjectiveVector(objective);
  DetailedCriteria oc = EffectiveOptimalityCriteria(params_.termination_criteria);
  oc.eps_optimal_dual_residual_absolute = kInf; oc.eps_optimal_dual_residual_relative = kInf;
  oc.eps_optimal_objective_gap_absolute = kInf; oc.eps_optimal_objective_gap_relative = kInf;
  p.termination_criteria.has_detailed = true; p.termination_criteria.has_simple = false;
  p.termination_criteria.detailed = oc;
  Vec y0;
  SetZero(sq().DualSharder(), y0);
  Solver sub(p, std::move(x_start), std::move(y0), step_size_, primal_weight_, pre_);
  SolveLog sub_log;
  timer_.Stop();
  SolverResult r = sub.Solve(IterationType::kPrimalFeasibility, interrupt, sub_log);
  timer_.Start();
  pre_->sq().SwapObjectiveVector(objective);
  log.feasibility_polishing_details.push_back(BuildFeasibilityPolishingDetails(POLISHING_PHASE_TYPE_PRIMAL_FEASIBILITY, iterations_completed_, p, r.solve_log));
  return r;
}

inline SolverResult Solver::TryDualPolishing(Vec y_start, int iteration_limit, const volatile int* interrupt, SolveLog& log) {  // :2616-2687
  Params p = params_;
  p.termination_criteria = ReduceWorkLimitsByPreviousWork(params_.termination_criteria, iteration_limit, TotalWorkSoFar(log));
  auto finite_to_zero = [](const Vec& in) { Vec out(in.size()); for (size_t i = 0; i < in.size(); ++i) out[i] = std::isfinite(in[i]) ? 0.0 : in[i]; return out; };
  Vec clb = finite_to_zero(qp().clb), cub = finite_to_zero(qp().cub), vlb = finite_to_zero(qp().vlb), vub = finite_to_zero(qp().vub);
  pre_->sq().SwapConstraintBounds(clb, cub);
  pre_->sq().SwapVariableBounds(vlb, vub);
  DetailedCriteria oc = EffectiveOptimalityCriteria(params_.termination_criteria);
  oc.eps_optimal_primal_residual_absolute = kInf; oc.eps_optimal_primal_residual_relative = kInf;
  oc.eps_optimal_objective_gap_absolute = kInf; oc.eps_optimal_objective_gap_relative = kInf;
  p.termination_criteria.has_detailed = true; p.termination_criteria.has_simple = false;
  p.termination_criteria.detailed = oc;
  Vec x0;
  SetZero(sq().PrimalSharder(), x0);
  Solver sub(p, std::move(x0), std::move(y_start), step_size_, primal_weight_, pre_);
  SolveLog sub_log;
  timer_.Stop();
  SolverResult r = sub.Solve(IterationType::kDualFeasibility, interrupt, sub_log);
  timer_.Start();
  pre_->sq().SwapConstraintBounds(clb, cub);
  pre_->sq().SwapVariableBounds(vlb, vub);
  log.feasibility_polishing_details.push_back(BuildFeasibilityPolishingDetails(POLISHING_PHASE_TYPE_DUAL_FEASIBILITY, iterations_completed_, p, r.solve_log));
  return r;
}

inline SolverResult Solver::Solve(IterationType type, const volatile int* interrupt, SolveLog log) {  // :2689-2774
  preprocessing_time_sec_ = log.preprocessing_time_sec;
  timer_.Start();
  x_start_ = CloneVector(x_, sq().PrimalSharder());
  y_start_ = CloneVector(y_, sq().DualSharder());
  ratio_last_two_step_sizes_ = 1;
  kty_ = TransposedMatVec(qp().K, y_, sq().KSharder());
  bool force_numerical = false;
  int next_polishing_iteration = 100;
  num_rejected_steps_ = 0;
  IterationStats polishing_work = WorkFromFeasibilityPolishing(log);
  for (iterations_completed_ = 0;; ++iterations_completed_) {
    auto maybe = MajorIterationAndTerminationCheck(type, force_numerical, interrupt, polishing_work, log);
    if (maybe.has_value()) return std::move(*maybe);
    if (params_.use_feasibility_polishing && type == IterationType::kNormal && iterations_completed_ >= next_polishing_iteration) {
      auto polished = TryFeasibilityPolishing(iterations_completed_ / 8, interrupt, log);
      if (polished.has_value()) return std::move(*polished);
      next_polishing_iteration *= 2;
      polishing_work = WorkFromFeasibilityPolishing(log);
    }
    InnerStepOutcome outcome;
    switch (params_.linesearch_rule) {
      case MALITSKY_POCK_LINESEARCH_RULE: outcome = TakeMalitskyPockStep(); break;
      case ADAPTIVE_LINESEARCH_RULE: outcome = TakeAdaptiveStep(); break;
      case CONSTANT_STEP_SIZE_RULE: outcome = TakeConstantSizeStep(); break;
      default: ORACLE_CHECK(false); outcome = InnerStepOutcome::kSuccessful;
    }
    if (outcome == InnerStepOutcome::kForceNumericalTermination) force_numerical = true;
  }
}

// Entry point (primal_dual_hybrid_gradient.cc:2787-2831).  CSC compression is a
// property of the caller's container and is checked at the C boundary.
inline SolverResult PrimalDualHybridGradient(Qp qp, const Params& params, const Vec* x0, const Vec* y0, const volatile int* interrupt = nullptr, Callback cb = nullptr) {
  const std::string perr = ValidateParams(params);
  if (!perr.empty()) return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, perr);
  const std::string derr = ValidateDimensions(qp);
  if (!derr.empty()) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "INVALID_ARGUMENT: " + derr);
  if (!HasValidBounds(qp)) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "The input problem has inconsistent bounds.");
  if (qp.objective_scaling_factor == 0) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "The objective scaling factor cannot be zero.");
  if (qp.q.has_value() && !qp.q->empty() && *std::min_element(qp.q->begin(), qp.q->end()) < 0.0)
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "The objective is not convex (i.e., the objective matrix contains negative entries).");
  if (params.use_feasibility_polishing && !qp.IsLp())
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, "use_feasibility_polishing is only implemented for linear programs.");
  if (params.use_glop)
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, "presolve_options.use_glop is not supported: Glop presolve is host-only and outside the restated path.");
  PreprocessSolver solver(std::move(qp), params);
  return solver.PreprocessAndSolve(params, x0, y0, interrupt, std::move(cb));
}

}  // namespace oracle
#endif  // PDLP_ORACLE_PDHG_IMPL_H_
