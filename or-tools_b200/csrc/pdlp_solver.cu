// Solver: the PDHG loop.  Host control flow of
//   ortools/pdlp/primal_dual_hybrid_gradient.cc:1773-2190 (stats, restarts, major
//   iterations), :2408-2687 (feasibility polishing), :2689-2774 (Solve)
// driving the device kernels of pdlp_step.cu.  The host only acts at "event" iterations
// (termination check, major iteration, artificial restart, work limit); between events
// the device runs trial steps back to back without a host round trip.
#include <algorithm>
#include <cstdio>
#include <chrono>
#include <cstdlib>
#include <thread>

#include "pdlp_driver.h"
#include "pdlp_logging.h"

namespace pdlp_b200 {

thread_local SolveCounters g_last_counters;
const SolveCounters& LastSolveCounters() { return g_last_counters; }

namespace {

constexpr double kInf = std::numeric_limits<double>::infinity();

double BoundGap(double lower, double upper) { return upper - lower; }

IterationStats AddWorkStats(IterationStats s, const IterationStats& extra) {  // pdhg.cc:2100-2113
  s.iteration_number = s.iteration_number + extra.iteration_number;
  s.cumulative_kkt_matrix_passes = s.cumulative_kkt_matrix_passes + extra.cumulative_kkt_matrix_passes;
  s.cumulative_rejected_steps = s.cumulative_rejected_steps + extra.cumulative_rejected_steps;
  s.cumulative_time_sec = s.cumulative_time_sec + extra.cumulative_time_sec;
  return s;
}

IterationStats WorkFromFeasibilityPolishing(const SolveLog& log) {  // pdhg.cc:2118-2126
  IterationStats r;
  r.iteration_number = 0; r.cumulative_kkt_matrix_passes = 0.0; r.cumulative_rejected_steps = 0; r.cumulative_time_sec = 0.0;
  for (const FeasibilityPolishingDetails& d : log.feasibility_polishing_details) r = AddWorkStats(std::move(r), d.solution_stats.get());
  return r;
}

bool TerminationReasonIsWorkLimit(int r) {  // pdhg.cc:2435-2440
  return r == TERMINATION_REASON_ITERATION_LIMIT || r == TERMINATION_REASON_TIME_LIMIT || r == TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT ||
         r == TERMINATION_REASON_INTERRUPTED_BY_USER;
}

}  // namespace

Solver::Solver(const PrimalDualHybridGradientParams& params, const double* x0, const double* y0, double initial_step_size, double initial_primal_weight,
               PreprocessSolver* preprocess_solver)
    : params_(params), pre_(preprocess_solver), ws_(preprocess_solver->ws()), problem_(preprocess_solver->problem()) {
  it_.Init(problem_.n, problem_.m);
  CopyVector(ws_, it_.x[0].get(), x0, problem_.n);
  CopyVector(ws_, it_.y[0].get(), y0, problem_.m);
  it_.avg_x.Zero(ws_.stream);
  it_.avg_y.Zero(ws_.stream);
  host_ = StepState{};
  host_.cur = 0;
  host_.step_size = initial_step_size;
  host_.primal_weight = initial_primal_weight;
  host_.ratio_last_two_step_sizes = 1.0;
  host_.kkt_pass_limit = kInf;
  host_.passes_per_rejected_step = params.linesearch_rule == MALITSKY_POCK_LINESEARCH_RULE ? 0.5 : 1.0;
  host_.reduction_exponent = params.adaptive_linesearch_parameters->step_size_reduction_exponent;
  host_.growth_exponent = params.adaptive_linesearch_parameters->step_size_growth_exponent;
  host_.mp_downscaling = params.malitsky_pock_parameters->step_size_downscaling_factor;
  host_.mp_contraction = params.malitsky_pock_parameters->linesearch_contraction_factor;
  host_.mp_interpolation = params.malitsky_pock_parameters->step_size_interpolation;
  host_.mp_tau = -1.0;
  host_.mp_new_tau = -1.0;
}

void Solver::PushState() {
  PDLP_CUDA_CHECK(cudaMemcpyAsync(it_.state.get(), &host_, sizeof(StepState), cudaMemcpyHostToDevice, ws_.stream));
  // host_ is pageable: the copy is staged before the call returns, so host_ may change after.
}
void Solver::PullState() {
  PDLP_CUDA_CHECK(cudaMemcpyAsync(&host_, it_.state.get(), sizeof(StepState), cudaMemcpyDeviceToHost, ws_.stream));
  ws_.Sync();
  if (host_.stopped_by_interrupt != 0) interrupts_seen_on_device_ += 1;
}

void Solver::EnableDeviceInterrupt(const InterruptFlag* interrupt) {
  interrupt_ = interrupt;
  if (interrupt == nullptr || pre_->sharded()) return;  // sharded: all ranks must stop at the same iteration, agreed on at host events
  if (mapped_interrupt_ == nullptr) {
    PDLP_CUDA_CHECK(cudaHostAlloc(reinterpret_cast<void**>(&mapped_interrupt_), sizeof(int), cudaHostAllocMapped));
    for (cudaEvent_t& e : chunk_done_) PDLP_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  *mapped_interrupt_ = 0;
  int* device_view = nullptr;
  PDLP_CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&device_view), mapped_interrupt_, 0));
  host_.interrupt_flag = device_view;
  if (ws_.stop_enqueue != mapped_interrupt_) outer_stop_enqueue_ = ws_.stop_enqueue;  // a polishing phase runs inside the main solver's Solve()
  ws_.stop_enqueue = mapped_interrupt_;
  // The solving thread spends a batch of trial steps inside kernel launches (the launch queue is shorter than a batch)
  // and then in a stream synchronize: a watcher forwards the caller's flag meanwhile.
  StopWatcher();
  watcher_stop_.store(false);
  watcher_ = std::thread([this] {
    while (!watcher_stop_.load(std::memory_order_relaxed)) {
      if (interrupt_->IsSet()) *reinterpret_cast<volatile int*>(mapped_interrupt_) = 1;
      std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
  });
}

void Solver::StopWatcher() {
  if (mapped_interrupt_ != nullptr && ws_.stop_enqueue == mapped_interrupt_) ws_.stop_enqueue = outer_stop_enqueue_;
  if (watcher_.joinable()) {
    watcher_stop_.store(true);
    watcher_.join();
  }
}

Solver::~Solver() {
  StopWatcher();
  if (mapped_interrupt_ != nullptr) {
    cudaStreamSynchronize(ws_.stream);  // no kernel may still read the flag
    cudaFreeHost(mapped_interrupt_);
  }
  for (cudaEvent_t e : chunk_done_) if (e != nullptr) cudaEventDestroy(e);
}

// Runs trial steps on the device until iterations_completed reaches the target, the
// KKT-pass limit is hit, or a step forces numerical termination.
void Solver::RunDeviceIterations(int target_iteration) {
  const TraceRange trace("pdlp/trial_steps");
  host_.target_iteration = target_iteration;
  host_.active = 1;
  host_.force_numerical = 0;
  host_.stopped_by_interrupt = 0;
  if (mapped_interrupt_ != nullptr) *reinterpret_cast<volatile int*>(mapped_interrupt_) = interrupt_->IsSet() ? 1 : 0;
  const TerminationCriteria& tc = params_.termination_criteria.get();
  // The work counted against kkt_matrix_pass_limit includes polishing work; the host
  // re-checks exactly at the next event, the device only needs a conservative stop.
  host_.kkt_pass_limit = tc.kkt_matrix_pass_limit - kkt_passes_from_polishing_;  // the reference tests full_work_stats (:2147), which includes polishing work
  PushState();
  const bool mp = params_.linesearch_rule == MALITSKY_POCK_LINESEARCH_RULE;
  const bool constant = params_.linesearch_rule == CONSTANT_STEP_SIZE_RULE;
  int guard = 0;
  for (;;) {
    const int need = target_iteration - host_.iterations_completed;
    const int trials = need + 2 + need / 8;
    auto enqueue = [&](int count) {
      if (mp) EnqueueMalitskyPockTrials(ws_, problem_, it_, count);
      else EnqueueAdaptiveTrials(ws_, problem_, it_, count, constant);
    };
    if (mapped_interrupt_ == nullptr) {
      enqueue(trials);
    } else {
      // With an interrupt flag the host must not run a whole batch ahead of the device: once the decision kernel has
      // switched the loop off, every launch still queued is a no-op that costs its launch latency (thousands of them
      // are tens of milliseconds).  At most two chunks are queued behind the one that is running.
      constexpr int kChunk = 32;
      int k = 0;
      for (int left = trials; left > 0; left -= kChunk, ++k) {
        enqueue(std::min(kChunk, left));
        PDLP_CUDA_CHECK(cudaEventRecord(chunk_done_[k & 1], ws_.stream));
        if (k >= 1) PDLP_CUDA_CHECK(cudaEventSynchronize(chunk_done_[(k - 1) & 1]));
        if (*reinterpret_cast<volatile int*>(mapped_interrupt_) != 0) break;
      }
    }
    PullState();
    if (host_.active == 0) break;
    if (++guard > 100000) throw std::runtime_error("pdlp_b200: device iteration loop made no progress");
  }
  // A delta exists once a step has been ACCEPTED (a forced numerical termination without an accepted step
  // advances the count but leaves no new delta).
  if (host_.iterations_completed - (host_.force_numerical != 0 && host_.accepted_last == 0 ? 1 : 0) > iterations_completed_) have_delta_ = true;
  iterations_completed_ = host_.iterations_completed;
}

// Smallest iteration count > k at which the host has something to do at the top of
// the loop (see MajorIterationAndTerminationCheck and Solve).
int Solver::NextEventIteration(int k, IterationType type, int next_polishing_iteration) const {
  const TerminationCriteria& tc = params_.termination_criteria.get();
  const int major = params_.major_iteration_frequency, check = params_.termination_check_frequency;
  if (params_.record_iteration_stats) return k + 1;
  const bool mp = params_.linesearch_rule == MALITSKY_POCK_LINESEARCH_RULE;
  int terms = host_.avg_num_terms;
  for (int e = k + 1;; ++e) {
    terms += 1;
    if (mp && !(host_.avg_weight_sum > 0.0) && e == k + 1) terms += 1;  // first M-P step adds the old x too
    const int cycle = e % major;
    if (cycle % check == 0) return e;                      // termination check (covers major iterations)
    if (e >= tc.iteration_limit) return e;
    if (params_.restart_strategy == ADAPTIVE_HEURISTIC && terms >= e / 2) return e;  // artificial restart, pdhg.cc:1915-1925
    if (params_.use_feasibility_polishing && type == IterationType::kNormal && e >= next_polishing_iteration) return e;
    if (e - k >= poll_iterations_ && (std::isfinite(tc.time_sec_limit.get()) || poll_always_)) return e;
  }
}

IterationStats Solver::CreateSimpleIterationStats(int restart_used) const {  // pdhg.cc:1773-1793
  IterationStats s;
  s.iteration_number = iterations_completed_;
  s.cumulative_rejected_steps = host_.num_rejected_steps;
  s.cumulative_kkt_matrix_passes = iterations_completed_ + host_.passes_per_rejected_step * host_.num_rejected_steps;
  s.cumulative_time_sec = Seconds();
  s.restart_used = restart_used;
  s.step_size = host_.step_size;
  s.primal_weight = host_.primal_weight;
  return s;
}

double Solver::DistanceTraveledFromLastStart(const double* x, const double* y) {  // pdhg.cc:1795-1804
  const double dp = SquaredDistance(ws_, x, it_.x_start.get(), problem_.n, Side::kSharded);
  const double dd = SquaredDistance(ws_, y, it_.y_start.get(), problem_.m, Side::kDual);
  return std::sqrt((0.5 * host_.primal_weight) * dp + (0.5 / host_.primal_weight) * dd);
}

Solver::Bounds Solver::BoundsAt(const double* x, const double* y, const double* kty_or_null, const double* kx_or_null) {
  // The radius (DistanceTraveledFromLastStart) comes out of the same passes as the gradients.
  int tr_passes = 0;
  const LocalizedBounds b = ComputeLocalizedLagrangianBounds(ws_, problem_, x, y, kty_or_null, host_.primal_weight, 0.0, params_.use_diagonal_qp_trust_region_solver,
                                                             params_.diagonal_qp_trust_region_solver_tolerance, &tr_passes, false, it_.x_start.get(), it_.y_start.get(),
                                                             kx_or_null);
  phase_.bounds += 1;
  phase_.tr_passes += tr_passes;
  return {b.lagrangian_value, b.lower_bound, b.upper_bound, b.radius};
}
Solver::Bounds Solver::BoundsAtCurrent() { return BoundsAt(X(), Y(), Kty(), event_.valid ? event_.kx_cur : nullptr); }  // pdhg.cc:1806-1816
Solver::Bounds Solver::BoundsAtAverage() {  // pdhg.cc:1818-1834
  const bool has_primal = host_.avg_weight_sum > 0.0, has_dual = host_.avg_weight_sum_dual > 0.0;
  // The event's products are those of (avg_x, avg_y); a side without weight falls back to the current iterate's.
  const double* kx = !event_.valid ? nullptr : (has_primal ? event_.kx_avg : event_.kx_cur);
  const double* kty = !event_.valid ? nullptr : (has_dual ? event_.kty_avg : Kty());
  const Bounds b = BoundsAt(PrimalAverage(), DualAverage(), kty, kx);
  // Without the event's products K^T y of the dual average is now in the bounds' scratch vector: a restart to the
  // average at this event copies it from there (or from the event's products).
  kty_of_average_at_ = (kty == nullptr && problem_.n > 0 && has_primal && has_dual) ? iterations_completed_ : -1;
  return b;
}

// One pass over each copy of K for every product this host event needs (one GPU; see EventProducts).
void Solver::PrepareEventProducts(const double* dx, const double* dy) {
  const TraceRange trace("pdlp/event_products");
  event_ = EventProducts{};
  if (pre_->sharded() || problem_.n == 0 || problem_.m == 0) return;
  const bool has_avg = host_.avg_weight_sum > 0.0 && host_.avg_weight_sum_dual > 0.0;
  if (event_m_[0].size() == 0) {
    for (DVec& v : event_m_) v.Resize(problem_.m);
    for (DVec& v : event_n_) v.Resize(problem_.n);
  }
  PointProducts points[3];
  int count = 0;
  points[count++] = PointProducts{X(), nullptr, event_m_[0].get(), nullptr};  // K^T y of the current iterate is cached
  event_.kx_cur = event_m_[0].get();
  if (has_avg) {
    points[count++] = PointProducts{it_.avg_x.get(), it_.avg_y.get(), event_m_[1].get(), event_n_[0].get()};
    event_.kx_avg = event_m_[1].get();
    event_.kty_avg = event_n_[0].get();
  }
  if (dx != nullptr && dy != nullptr) {
    points[count++] = PointProducts{dx, dy, event_m_[2].get(), event_n_[1].get()};
    event_.kx_diff = event_m_[2].get();
    event_.kty_diff = event_n_[1].get();
  }
  ComputeProducts(ws_, problem_, points, count);
  event_.valid = true;
}

bool Solver::ShouldDoAdaptiveRestartHeuristic(double candidate_normalized_gap) const {  // pdhg.cc:1853-1867
  const double ratio = candidate_normalized_gap / normalized_gap_at_last_restart_;
  if (ratio < params_.sufficient_reduction_for_restart) return true;
  if (ratio < params_.necessary_reduction_for_restart && candidate_normalized_gap > normalized_gap_at_last_trial_) return true;
  return false;
}

namespace {
bool AverageHasBetterPotential(double avg_gap, double avg_radius, double cur_gap, double cur_radius) {  // pdhg.cc:1836-1843
  return avg_gap / (avg_radius * avg_radius) < cur_gap / (cur_radius * cur_radius);
}
}  // namespace

int Solver::DetermineDistanceBasedRestartChoice() {  // pdhg.cc:1869-1902
  if (host_.avg_num_terms == 0) return RESTART_CHOICE_NO_RESTART;
  if (distance_info_.length_of_last_restart_period == 0) return RESTART_CHOICE_RESTART_TO_AVERAGE;
  const int restart_period_length = host_.avg_num_terms;
  const double distance_moved_this_restart_period_by_average = DistanceTraveledFromLastStart(it_.avg_x.get(), it_.avg_y.get());
  const double distance_moved_last_restart_period = distance_info_.distance_moved_last_restart_period;
  if ((distance_moved_this_restart_period_by_average / restart_period_length) <
      params_.sufficient_reduction_for_restart * (distance_moved_last_restart_period / distance_info_.length_of_last_restart_period)) {
    const Bounds a = BoundsAtAverage(), c = BoundsAtCurrent();
    return AverageHasBetterPotential(BoundGap(a.lower_bound, a.upper_bound), a.radius, BoundGap(c.lower_bound, c.upper_bound), c.radius)
               ? RESTART_CHOICE_RESTART_TO_AVERAGE
               : RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
  }
  return RESTART_CHOICE_NO_RESTART;
}

int Solver::ChooseRestartToApply(bool is_major_iteration) {  // pdhg.cc:1904-1965
  const TraceRange trace("pdlp/restart_choice");
  if (!(host_.avg_weight_sum > 0.0) && !(host_.avg_weight_sum_dual > 0.0)) return RESTART_CHOICE_NO_RESTART;
  const int restart_length = host_.avg_num_terms;
  if (restart_length >= iterations_completed_ / 2 && params_.restart_strategy == ADAPTIVE_HEURISTIC) {
    const Bounds a = BoundsAtAverage(), c = BoundsAtCurrent();
    return AverageHasBetterPotential(BoundGap(a.lower_bound, a.upper_bound), a.radius, BoundGap(c.lower_bound, c.upper_bound), c.radius)
               ? RESTART_CHOICE_RESTART_TO_AVERAGE
               : RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
  }
  if (!is_major_iteration) return RESTART_CHOICE_NO_RESTART;
  switch (static_cast<int>(params_.restart_strategy)) {
    case NO_RESTARTS: return RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
    case EVERY_MAJOR_ITERATION: return RESTART_CHOICE_RESTART_TO_AVERAGE;
    case ADAPTIVE_HEURISTIC: {
      const Bounds a = BoundsAtAverage(), c = BoundsAtCurrent();
      double normalized_gap;
      int choice;
      if (AverageHasBetterPotential(BoundGap(a.lower_bound, a.upper_bound), a.radius, BoundGap(c.lower_bound, c.upper_bound), c.radius)) {
        normalized_gap = BoundGap(a.lower_bound, a.upper_bound) / a.radius;
        choice = RESTART_CHOICE_RESTART_TO_AVERAGE;
      } else {
        normalized_gap = BoundGap(c.lower_bound, c.upper_bound) / c.radius;
        choice = RESTART_CHOICE_WEIGHTED_AVERAGE_RESET;
      }
      if (ShouldDoAdaptiveRestartHeuristic(normalized_gap)) return choice;
      normalized_gap_at_last_trial_ = normalized_gap;
      return RESTART_CHOICE_NO_RESTART;
    }
    case ADAPTIVE_DISTANCE_BASED: return DetermineDistanceBasedRestartChoice();
    default: throw std::logic_error("pdlp_b200: restart strategy not specified");
  }
}

double Solver::ComputeNewPrimalWeight() {  // pdhg.cc:1983-2010
  const double primal_distance = std::sqrt(SquaredDistance(ws_, X(), it_.x_start.get(), problem_.n, Side::kSharded));
  const double dual_distance = std::sqrt(SquaredDistance(ws_, Y(), it_.y_start.get(), problem_.m, Side::kDual));
  constexpr double kNonzeroTol = 1.0e-10;
  if (primal_distance <= kNonzeroTol || primal_distance >= 1.0 / kNonzeroTol || dual_distance <= kNonzeroTol || dual_distance >= 1.0 / kNonzeroTol) {
    return host_.primal_weight;
  }
  const double smoothing = params_.primal_weight_update_smoothing;
  const double new_primal_weight = std::exp(smoothing * std::log(dual_distance / primal_distance) + (1.0 - smoothing) * std::log(host_.primal_weight));
  if (pre_->Logs(params_, 4)) LogInfoWithoutPrefix(Format("New computed primal weight is %g at iteration %d", new_primal_weight, iterations_completed_));  // :2004-2008
  return new_primal_weight;
}

void Solver::ApplyRestartChoice(int restart_to_apply) {  // pdhg.cc:2042-2094
  const TraceRange trace("pdlp/apply_restart");
  switch (restart_to_apply) {
    case RESTART_CHOICE_UNSPECIFIED:
    case RESTART_CHOICE_NO_RESTART:
      return;
    case RESTART_CHOICE_WEIGHTED_AVERAGE_RESET:
      if (pre_->Logs(params_, 4)) LogInfoWithoutPrefix(Format("Restarted to current on iteration %d after %d iterations", iterations_completed_, host_.avg_num_terms));  // :2048-2052
      break;
    case RESTART_CHOICE_RESTART_TO_AVERAGE:
      if (pre_->Logs(params_, 4)) LogInfoWithoutPrefix(Format("Restarted to average on iteration %d after %d iterations", iterations_completed_, host_.avg_num_terms));  // :2055-2059
      CopyVector(ws_, it_.x[host_.cur].get(), it_.avg_x.get(), problem_.n);
      CopyVector(ws_, it_.y[host_.cur].get(), it_.avg_y.get(), problem_.m);
      // K^T y of the new current iterate: the same product, by the same kernel over the same values, that the restart
      // choice of this event left in the bounds' scratch vector (nothing between here and there writes scratch_n[1]).
      if (event_.valid && event_.kty_avg != nullptr) {
        CopyVector(ws_, it_.kty[host_.cur].get(), event_.kty_avg, problem_.n);
        event_.kx_cur = event_.kx_avg;  // the current iterate IS the average now: the bounds below reuse its K x
      } else {
        event_.valid = false;
        if (kty_of_average_at_ == iterations_completed_) CopyVector(ws_, it_.kty[host_.cur].get(), ws_.scratch_n[1].get(), problem_.n);
        else RowDot(ws_, problem_.K.by_col, it_.y[host_.cur].get(), it_.kty[host_.cur].get());
      }
      break;
  }
  kty_of_average_at_ = -1;
  host_.primal_weight = ComputeNewPrimalWeight();
  host_.ratio_last_two_step_sizes = 1.0;
  if (params_.restart_strategy == ADAPTIVE_HEURISTIC) {
    const Bounds b = BoundsAtCurrent();
    normalized_gap_at_last_restart_ = BoundGap(b.lower_bound, b.upper_bound) / b.radius;
    normalized_gap_at_last_trial_ = kInf;
  } else if (params_.restart_strategy == ADAPTIVE_DISTANCE_BASED) {
    distance_info_.distance_moved_last_restart_period = DistanceTraveledFromLastStart(X(), Y());
    distance_info_.length_of_last_restart_period = host_.avg_num_terms;
  }
  it_.avg_x.Zero(ws_.stream);
  it_.avg_y.Zero(ws_.stream);
  host_.avg_weight_sum = 0.0;
  host_.avg_weight_sum_dual = 0.0;
  host_.avg_num_terms = 0;
  CopyVector(ws_, it_.x_start.get(), X(), problem_.n);
  CopyVector(ws_, it_.y_start.get(), Y(), problem_.m);
}

Solver::Outcome Solver::PickSolutionAndConstructOutcome(const IterationStats& stats, int reason, int point_type, SolveLog log) {  // pdhg.cc:2012-2040, 303-316
  Outcome out;
  switch (point_type) {
    case POINT_TYPE_CURRENT_ITERATE:
      out.x = X(); out.y = Y();
      break;
    case POINT_TYPE_ITERATE_DIFFERENCE: {
      double* dx = ws_.scratch_n[2].get();
      double* dy = ws_.scratch_m[2].get();
      SubtractVectors(ws_, dx, it_.x[host_.cur].get(), it_.x[host_.cur ^ 1].get(), problem_.n);
      SubtractVectors(ws_, dy, it_.y[host_.cur].get(), it_.y[host_.cur ^ 1].get(), problem_.m);
      out.x = dx; out.y = dy;
      break;
    }
    case POINT_TYPE_AVERAGE_ITERATE:
    case POINT_TYPE_PRESOLVER_SOLUTION:
      out.x = PrimalAverage(); out.y = DualAverage();
      break;
    default:
      point_type = POINT_TYPE_AVERAGE_ITERATE;
      out.x = PrimalAverage(); out.y = DualAverage();
      break;
  }
  log.iteration_count = stats.iteration_number;
  log.termination_reason = reason;
  log.solution_type = point_type;
  log.solve_time_sec = stats.cumulative_time_sec;
  log.solution_stats.mut() = stats;
  out.log = std::move(log);
  return out;
}

std::optional<Solver::Outcome> Solver::MajorIterationAndTerminationCheck(IterationType type, bool force_numerical_termination,
                                                                         const InterruptFlag* interrupt,
                                                                         const IterationStats& work_from_feasibility_polishing, SolveLog& log) {  // pdhg.cc:2128-2190
  if (pre_->sharded()) {
    AgreeOnClockAndInterrupt(interrupt);
    interrupt = &agreed_interrupt_;
  }
  const int major_iteration_cycle = iterations_completed_ % params_.major_iteration_frequency;
  const bool is_major_iteration = major_iteration_cycle == 0 && iterations_completed_ > 0;
  const double t_begin = timer_.Get();
  // What this event will multiply by K: the termination check its three points, the restart choice the current iterate
  // and the average.  All of it comes from one pass over each matrix copy, made first (one GPU).  The predicates below
  // only decide what is computed ahead; a product that turns out to be missing is computed where it is needed.
  const double* dx = nullptr;
  const double* dy = nullptr;
  {
    const IterationStats preview = AddWorkStats(CreateSimpleIterationStats(RESTART_CHOICE_NO_RESTART), work_from_feasibility_polishing);
    const bool will_check = major_iteration_cycle % params_.termination_check_frequency == 0 || force_numerical_termination ||
                            CheckSimpleTerminationCriteria(params_.termination_criteria.get(), preview, interrupt).has_value();
    const bool has_weight = host_.avg_weight_sum > 0.0 || host_.avg_weight_sum_dual > 0.0;
    const bool will_bound = !force_numerical_termination && has_weight && params_.restart_strategy == ADAPTIVE_HEURISTIC &&
                            (is_major_iteration || host_.avg_num_terms >= iterations_completed_ / 2);
    if (will_check && have_delta_) {
      double* dxb = ws_.scratch_n[2].get();
      double* dyb = ws_.scratch_m[2].get();
      SubtractVectors(ws_, dxb, it_.x[host_.cur].get(), it_.x[host_.cur ^ 1].get(), problem_.n);
      SubtractVectors(ws_, dyb, it_.y[host_.cur].get(), it_.y[host_.cur ^ 1].get(), problem_.m);
      dx = dxb; dy = dyb;
    }
    if (will_check || will_bound) PrepareEventProducts(dx, dy);
    else event_ = EventProducts{};
  }
  const int restart = force_numerical_termination ? static_cast<int>(RESTART_CHOICE_NO_RESTART) : ChooseRestartToApply(is_major_iteration);
  const double t_chosen = timer_.Get();
  phase_.events += 1;
  phase_.restart_choice += t_chosen - t_begin;
  IterationStats stats = CreateSimpleIterationStats(restart);
  const IterationStats full_work_stats = AddWorkStats(stats, work_from_feasibility_polishing);
  const bool check_termination = major_iteration_cycle % params_.termination_check_frequency == 0 ||
                                 CheckSimpleTerminationCriteria(params_.termination_criteria.get(), full_work_stats, interrupt).has_value() ||
                                 force_numerical_termination;
  if (check_termination) {
    const bool has_avg = host_.avg_weight_sum > 0.0 && host_.avg_weight_sum_dual > 0.0;
    if (have_delta_ && dx == nullptr) {  // a limit was reached between the preview and here
      double* dxb = ws_.scratch_n[2].get();
      double* dyb = ws_.scratch_m[2].get();
      SubtractVectors(ws_, dxb, it_.x[host_.cur].get(), it_.x[host_.cur ^ 1].get(), problem_.n);
      SubtractVectors(ws_, dyb, it_.y[host_.cur].get(), it_.y[host_.cur ^ 1].get(), problem_.m);
      dx = dxb; dy = dyb;
    }
    PreprocessSolver::ProductsOfCheck products;
    if (event_.valid) {
      products.current = {event_.kx_cur, Kty()};
      if (has_avg) products.average = {event_.kx_avg, event_.kty_avg};
      products.difference = {event_.kx_diff, event_.kty_diff};
    }
    const TraceRange trace("pdlp/termination_check");
    const auto maybe_reason = pre_->UpdateIterationStatsAndCheckTermination(
        params_, force_numerical_termination, X(), Y(), has_avg ? it_.avg_x.get() : nullptr, has_avg ? it_.avg_y.get() : nullptr, dx, dy,
        it_.x_start.get(), it_.y_start.get(), interrupt, type, full_work_stats, stats, &products);
    if (params_.record_iteration_stats) log.iteration_stats.push_back(stats);
    phase_.checks += 1;
    phase_.termination += timer_.Get() - t_chosen;
    if (maybe_reason.has_value()) {
      const IterationStats terminating_full_stats = AddWorkStats(stats, work_from_feasibility_polishing);
      return PickSolutionAndConstructOutcome(terminating_full_stats, maybe_reason->reason, maybe_reason->type, std::move(log));
    }
  } else if (params_.record_iteration_stats) {
    log.iteration_stats.push_back(stats);
  }
  const double t_checked = timer_.Get();
  ApplyRestartChoice(restart);
  if (restart == RESTART_CHOICE_WEIGHTED_AVERAGE_RESET || restart == RESTART_CHOICE_RESTART_TO_AVERAGE) {
    ws_.Sync();
    phase_.restarts += 1;
    phase_.apply_restart += timer_.Get() - t_checked;
  }
  event_ = EventProducts{};  // the iterate moves on
  return std::nullopt;
}

void Solver::AgreeOnClockAndInterrupt(const InterruptFlag* interrupt) {
  double v[2] = {preprocessing_time_sec_ + timer_.Get(), (interrupt != nullptr && interrupt->IsSet()) ? 1.0 : 0.0};
  pre_->AgreeOnMax(v, 2);
  const int done = iterations_completed_ - agreed_seconds_iteration_;
  if (done > 0 && v[0] > agreed_seconds_) {
    const double per_iteration = (v[0] - agreed_seconds_) / done;
    poll_iterations_ = static_cast<int>(std::min(4096.0, std::max(1.0, 0.02 / per_iteration)));
  }
  agreed_seconds_ = std::max(agreed_seconds_, v[0]);
  agreed_seconds_iteration_ = iterations_completed_;
  agreed_interrupt_.agreed = v[1] != 0.0 ? 1 : 0;
}

IterationStats Solver::TotalWorkSoFar(const SolveLog& log) const {  // pdhg.cc:2408-2413
  return AddWorkStats(CreateSimpleIterationStats(RESTART_CHOICE_NO_RESTART), WorkFromFeasibilityPolishing(log));
}

// pdhg.cc:2689-2774.
Solver::Outcome Solver::Solve(IterationType type, const InterruptFlag* interrupt, SolveLog log) {
  preprocessing_time_sec_ = log.preprocessing_time_sec;
  timer_.Start();
  EnableDeviceInterrupt(interrupt);
  // One GPU: the decision kernel reads the flag itself, no host event is needed for it.  Sharded: the host polls.
  poll_always_ = interrupt != nullptr && mapped_interrupt_ == nullptr;
  if (pre_->sharded()) {
    double v = poll_always_ ? 1.0 : 0.0;
    pre_->AgreeOnMax(&v, 1);
    poll_always_ = v != 0.0;
    agreed_seconds_ = 0.0;
    agreed_seconds_iteration_ = 0;
  }
  CopyVector(ws_, it_.x_start.get(), X(), problem_.n);
  CopyVector(ws_, it_.y_start.get(), Y(), problem_.m);
  host_.ratio_last_two_step_sizes = 1.0;
  RowDot(ws_, problem_.K.by_col, Y(), it_.kty[host_.cur].get());
  bool force_numerical_termination = false;
  int next_feasibility_polishing_iteration = 100;
  host_.num_rejected_steps = 0;
  host_.iterations_completed = 0;
  iterations_completed_ = 0;
  IterationStats work_from_feasibility_polishing = WorkFromFeasibilityPolishing(log);
  kkt_passes_from_polishing_ = work_from_feasibility_polishing.cumulative_kkt_matrix_passes;
  struct EventPair {  // destroyed on every way out of Solve(), a throwing kernel launch or collective included
    cudaEvent_t begin = nullptr, end = nullptr;
    EventPair() { PDLP_CUDA_CHECK(cudaEventCreate(&begin)); PDLP_CUDA_CHECK(cudaEventCreate(&end)); }
    ~EventPair() { if (begin != nullptr) cudaEventDestroy(begin); if (end != nullptr) cudaEventDestroy(end); }
    EventPair(const EventPair&) = delete;
    EventPair& operator=(const EventPair&) = delete;
  } events;
  PDLP_CUDA_CHECK(cudaEventRecord(events.begin, ws_.stream));
  auto finish = [&](Outcome o) {
    // A sticky error raised inside the loop surfaces here at the latest, not in the caller's next CUDA call.
    PDLP_CUDA_CHECK(cudaEventRecord(events.end, ws_.stream));
    PDLP_CUDA_CHECK(cudaEventSynchronize(events.end));
    float ms = 0.0f;
    PDLP_CUDA_CHECK(cudaEventElapsedTime(&ms, events.begin, events.end));
    device_loop_seconds_ = ms * 1e-3;
    StopWatcher();
    if (std::getenv("PDLP_B200_DEBUG_TIMING") != nullptr) {
      std::fprintf(stderr, "pdlp_b200 loop phases: %d iterations, %lld trials, steps %.4f s; %d events: restart choice %.4f s (%d bounds, %d trust-region passes), "
                   "%d termination checks %.4f s, %d restarts applied %.4f s; device loop %.4f s\n", iterations_completed_, static_cast<long long>(host_.total_trials),
                   phase_.steps, phase_.events, phase_.restart_choice, phase_.bounds, phase_.tr_passes, phase_.checks, phase_.termination, phase_.restarts,
                   phase_.apply_restart, device_loop_seconds_);
    }
    return o;
  };
  for (;;) {
    auto maybe_result = MajorIterationAndTerminationCheck(type, force_numerical_termination, interrupt, work_from_feasibility_polishing, log);
    if (maybe_result.has_value()) return finish(std::move(*maybe_result));
    if (params_.use_feasibility_polishing && type == IterationType::kNormal && iterations_completed_ >= next_feasibility_polishing_iteration) {
      auto polished = TryFeasibilityPolishing(iterations_completed_ / 8, interrupt, log);
      if (polished.has_value()) return finish(std::move(*polished));
      next_feasibility_polishing_iteration *= 2;
      work_from_feasibility_polishing = WorkFromFeasibilityPolishing(log);
      kkt_passes_from_polishing_ = work_from_feasibility_polishing.cumulative_kkt_matrix_passes;
    }
    const int target = NextEventIteration(iterations_completed_, type, next_feasibility_polishing_iteration);
    const double t0 = timer_.Get();
    const int before = iterations_completed_;
    RunDeviceIterations(target);
    // Keep the polling interval near 20 ms of work when a time limit or interrupt is live.
    const double dt = timer_.Get() - t0;
    phase_.steps += dt;
    const int done = iterations_completed_ - before;
    if (done > 0 && dt > 0.0 && !pre_->sharded()) {
      const double per_iteration = dt / done;
      poll_iterations_ = static_cast<int>(std::min(4096.0, std::max(1.0, 0.02 / per_iteration)));
    }
    if (host_.force_numerical != 0) {
      if (!force_numerical_termination && pre_->Logs(params_, 2)) LogInfoWithoutPrefix(Format("Forced numerical termination at iteration %d", iterations_completed_));  // :2199-2204
      force_numerical_termination = true;
    }
  }
}

// ---- feasibility polishing (pdhg.cc:2442-2687) ----
namespace {
TerminationCriteria ReduceWorkLimitsByPreviousWork(TerminationCriteria c, int iteration_limit, const IterationStats& previous_work) {  // :2554-2566
  c.iteration_limit = std::max(0, std::min(iteration_limit, c.iteration_limit - previous_work.iteration_number));
  c.kkt_matrix_pass_limit = std::max(0.0, c.kkt_matrix_pass_limit - previous_work.cumulative_kkt_matrix_passes);
  c.time_sec_limit = std::max(0.0, c.time_sec_limit - previous_work.cumulative_time_sec);
  return c;
}
FeasibilityPolishingDetails BuildFeasibilityPolishingDetails(int phase, int main_iteration_count, const PrimalDualHybridGradientParams& params,
                                                             const SolveLog& log) {  // :2415-2431
  FeasibilityPolishingDetails d;
  d.polishing_phase_type = phase;
  d.main_iteration_count = main_iteration_count;
  d.params.mut() = params;
  d.termination_reason = static_cast<int>(log.termination_reason);
  d.iteration_count = log.iteration_count;
  d.solve_time_sec = log.solve_time_sec;
  d.solution_stats.mut() = log.solution_stats.get();
  d.solution_type = static_cast<int>(log.solution_type);
  d.iteration_stats = log.iteration_stats;
  return d;
}
struct FiniteToZeroHost {};
}  // namespace

std::optional<Solver::Outcome> Solver::TryFeasibilityPolishing(int iteration_limit, const InterruptFlag* interrupt, SolveLog& log) {
  const DetailedOptimalityCriteria optimality_criteria = EffectiveOptimalityCriteria(params_.termination_criteria.get());
  const double* average_primal = PrimalAverage();
  const double* average_dual = DualAverage();
  ConvergenceInformation first_convergence_info;
  pre_->ComputeConvergenceAndInfeasibility(params_, average_primal, average_dual, it_.x_start.get(), it_.y_start.get(), POINT_TYPE_AVERAGE_ITERATE,
                                           &first_convergence_info, nullptr, nullptr);
  if (!ObjectiveGapMet(optimality_criteria, first_convergence_info)) {
    if (pre_->Logs(params_, 2)) LogInfoWithoutPrefix("Skipping feasibility polishing because the objective gap is too large.");  // :2457-2461
    return std::nullopt;
  }

  const int64_t n = problem_.n, m = problem_.m;
  auto run_phase = [&](bool primal_phase, DVec* result_holder) -> int {
    PrimalDualHybridGradientParams phase_params = params_;
    phase_params.termination_criteria.mut() = ReduceWorkLimitsByPreviousWork(params_.termination_criteria.get(), iteration_limit, TotalWorkSoFar(log));
    DetailedOptimalityCriteria oc = EffectiveOptimalityCriteria(params_.termination_criteria.get());
    if (primal_phase) {
      oc.eps_optimal_dual_residual_absolute = kInf; oc.eps_optimal_dual_residual_relative = kInf;
    } else {
      oc.eps_optimal_primal_residual_absolute = kInf; oc.eps_optimal_primal_residual_relative = kInf;
    }
    oc.eps_optimal_objective_gap_absolute = kInf; oc.eps_optimal_objective_gap_relative = kInf;
    TerminationCriteria& ptc = phase_params.termination_criteria.mut();
    ptc.simple_optimality_criteria.has = false;
    ptc.eps_optimal_absolute.has = false;
    ptc.eps_optimal_relative.has = false;
    ptc.detailed_optimality_criteria.mut() = oc;

    // Swap in the phase's problem data (pdhg.cc:2568-2687): zero objective for the primal
    // phase; finite bounds mapped to zero for the dual phase.
    DVec zero_c, h_clb, h_cub, h_vlb, h_vub;
    if (primal_phase) {
      zero_c.Resize(n > 0 ? n : 1);
      zero_c.Zero(ws_.stream);
      problem_.c.swap(zero_c);
    } else {
      auto homogeneous = [&](const DVec& src, int64_t count, DVec* dst) {
        std::vector<double> h(count);
        src.Download(h.data(), count, ws_.stream);
        ws_.Sync();
        for (double& v : h) if (std::isfinite(v)) v = 0.0;
        dst->Resize(count > 0 ? count : 1);
        dst->Upload(h.data(), count, ws_.stream);
        ws_.Sync();
      };
      homogeneous(problem_.clb, m, &h_clb); homogeneous(problem_.cub, m, &h_cub);
      homogeneous(problem_.vlb, n, &h_vlb); homogeneous(problem_.vub, n, &h_vub);
      problem_.clb.swap(h_clb); problem_.cub.swap(h_cub); problem_.vlb.swap(h_vlb); problem_.vub.swap(h_vub);
    }
    DVec zeros(primal_phase ? (m > 0 ? m : 1) : (n > 0 ? n : 1));
    zeros.Zero(ws_.stream);
    const double* start_x = primal_phase ? average_primal : zeros.get();
    const double* start_y = primal_phase ? zeros.get() : average_dual;
    int reason;
    if (pre_->Logs(params_, 2)) LogInfoWithoutPrefix(primal_phase ? "Starting primal feasibility polishing" : "Starting dual feasibility polishing");  // :2465, :2491
    {
      Solver sub(phase_params, start_x, start_y, host_.step_size, host_.primal_weight, pre_);
      SolveLog sub_log;
      timer_.Stop();
      Outcome r = sub.Solve(primal_phase ? IterationType::kPrimalFeasibility : IterationType::kDualFeasibility, interrupt, sub_log);
      timer_.Start();
      host_.total_trials += sub.TrialSteps();
      reason = r.log.termination_reason;
      if (pre_->Logs(params_, 2)) {  // :2471-2475, :2497-2501
        LogInfoWithoutPrefix(std::string(primal_phase ? "Primal" : "Dual") + " feasibility polishing termination reason: " + TerminationReasonName(reason));
      }
      result_holder->Resize(primal_phase ? (n > 0 ? n : 1) : (m > 0 ? m : 1));
      CopyVector(ws_, result_holder->get(), primal_phase ? r.x : r.y, primal_phase ? n : m);
      ws_.Sync();
      log.feasibility_polishing_details.push_back(BuildFeasibilityPolishingDetails(
          primal_phase ? POLISHING_PHASE_TYPE_PRIMAL_FEASIBILITY : POLISHING_PHASE_TYPE_DUAL_FEASIBILITY, iterations_completed_, phase_params, r.log));
    }
    if (primal_phase) {
      problem_.c.swap(zero_c);
    } else {
      problem_.clb.swap(h_clb); problem_.cub.swap(h_cub); problem_.vlb.swap(h_vlb); problem_.vub.swap(h_vub);
    }
    return reason;
  };

  const int primal_reason = run_phase(true, &polished_x_);
  if (TerminationReasonIsWorkLimit(primal_reason)) return std::nullopt;
  if (primal_reason != TERMINATION_REASON_OPTIMAL) return std::nullopt;
  const int dual_reason = run_phase(false, &polished_y_);
  if (TerminationReasonIsWorkLimit(dual_reason)) return std::nullopt;
  if (dual_reason != TERMINATION_REASON_OPTIMAL) return std::nullopt;

  IterationStats full_stats = TotalWorkSoFar(log);
  ConvergenceInformation ci;
  pre_->ComputeConvergenceAndInfeasibility(params_, polished_x_.get(), polished_y_.get(), it_.x_start.get(), it_.y_start.get(),
                                           POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION, &ci, nullptr, nullptr);
  full_stats.convergence_information.push_back(ci);
  const auto earned = CheckIterateTerminationCriteria(params_.termination_criteria.get(), full_stats, pre_->OriginalBoundNorms(), false);
  if (earned.has_value()) {
    Outcome out;
    out.x = polished_x_.get();
    out.y = polished_y_.get();
    SolveLog l = log;
    l.iteration_count = full_stats.iteration_number;
    l.termination_reason = earned->reason;
    l.solution_type = POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION;
    l.solve_time_sec = full_stats.cumulative_time_sec;
    l.solution_stats.mut() = full_stats;
    out.log = std::move(l);
    return out;
  }
  return std::nullopt;
}

}  // namespace pdlp_b200
