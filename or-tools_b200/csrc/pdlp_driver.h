// Internal declarations of the host driver: the two classes that mirror
// PreprocessSolver and Solver of ortools/pdlp/primal_dual_hybrid_gradient.cc:318-709.
// The control flow stays on the host; vectors never leave the device until the result
// is built.
#ifndef PDLP_B200_DRIVER_H_
#define PDLP_B200_DRIVER_H_

#include <chrono>
#include <atomic>
#include <memory>
#include <thread>

#include "pdlp_ops.cuh"
#include "primal_dual_hybrid_gradient.h"

namespace pdlp_b200 {

// Borrowed view of a QuadraticProgram (the C ABI passes caller memory straight through).
struct QpView {
  int64_t n = 0, m = 0, nnz = 0;
  const int64_t* col_ptr = nullptr;
  const int64_t* row_idx = nullptr;
  const int32_t* col_ptr32 = nullptr;  // when set, the 32-bit pair replaces col_ptr / row_idx
  const int32_t* row_idx32 = nullptr;
  const double* val = nullptr;
  const double* c = nullptr;
  const double* q = nullptr;  // null for an LP
  const double* vlb = nullptr; const double* vub = nullptr;
  const double* clb = nullptr; const double* cub = nullptr;
  double objective_offset = 0.0, objective_scaling_factor = 1.0;
  const char* name = nullptr;
  // Column-sharded solve: this view is the block of columns
  // [col_offset, col_offset + n) of a problem with global_n columns; rank / world / unique_id
  // identify the NCCL communicator.  world == 1: the whole problem, no communicator.
  // m is then this rank's row count: rows [0, num_shared_rows) are shared by all ranks in the same
  // order, the rest touch this rank's columns only (num_shared_rows < 0: all m rows are shared).
  int64_t global_n = -1, col_offset = 0, num_shared_rows = -1;
  int rank = 0, world = 1;
  const void* nccl_unique_id = nullptr;  // 128 bytes
  int64_t GlobalN() const { return global_n >= 0 ? global_n : n; }
};

using IterationCallback = std::function<void(const IterationCallbackInfo&)>;

// Caller-owned destinations for the solution (the C ABI's output arrays): when given, the three
// vectors are copied from the device straight into them and SolverResult's own vectors stay empty,
// instead of device -> SolverResult -> caller.  A null member is skipped.
struct SolutionSink {
  double* primal = nullptr;
  double* dual = nullptr;
  double* reduced_costs = nullptr;
  bool filled = false;
};

class WallTimer {
 public:
  void Start() { running_ = true; t0_ = Clock::now(); }
  void Stop() { if (running_) { acc_ += Elapsed(); running_ = false; } }
  double Get() const { return acc_ + (running_ ? Elapsed() : 0.0); }
 private:
  using Clock = std::chrono::steady_clock;
  double Elapsed() const { return std::chrono::duration<double>(Clock::now() - t0_).count(); }
  Clock::time_point t0_;
  double acc_ = 0.0;
  bool running_ = false;
};

SolverResult ErrorSolverResult(int reason, const std::string& message);  // pdhg.cc:717-725
extern thread_local SolveCounters g_last_counters;

class Solver;

// K x / K^T y of the three points of a termination check when the caller already has them (Solver computes the
// products of all points of a host event with one stream of each matrix copy); a null entry is computed where needed.
struct ProductsOfPoint { const double* kx = nullptr; const double* kty = nullptr; };
struct ProductsOfCheck { ProductsOfPoint current, average, difference; };

class PreprocessSolver;

// What the reference keeps in PresolveInfo (pdhg.cc:395-417) when Glop presolve ran: the original problem
// (resident on the device through a PreprocessSolver of its own, never rescaled) and the way back from a
// solution of the presolved problem.  Here the presolver is the caller's (Glop is host code outside this
// library): `postsolve` maps UNSCALED presolved (x, y) host vectors to original-space host vectors, the
// role of glop::MainLpPreprocessor::RecoverSolution in RecoverOriginalSolution (:1288-1341).
struct PresolveHook {
  PreprocessSolver* original = nullptr;
  std::function<void(const double* presolved_x, const double* presolved_y, double* original_x, double* original_y)> postsolve;
  DVec x, y;  // the recovered point on the device (original sizes)
};

// pdhg.cc:318-502.
class PreprocessSolver {
 public:
  PreprocessSolver(const QpView& qp, const PrimalDualHybridGradientParams& params);
  ~PreprocessSolver();
  SolverResult PreprocessAndSolve(const PrimalDualHybridGradientParams& params, const double* x0, const double* y0,
                                  const InterruptFlag* interrupt, IterationCallback callback);
  void set_solution_sink(SolutionSink* sink) { sink_ = sink; }
  // The solve runs on this (presolved) problem; statistics, termination tests and the result are those of
  // hook->original at the recovered point (pdhg.cc:1530-1548, :1571-1585).
  void set_presolve(PresolveHook* hook) { presolve_ = hook; }
  // The presolver solved the problem by itself (pdhg.cc:986-1034): no iterations, the result is the recovered
  // empty solution with `presolve_reason` (a TerminationReason) unless the point earns a better one.
  SolverResult FinishPresolvedOnly(const PrimalDualHybridGradientParams& params, int presolve_reason, const InterruptFlag* interrupt, IterationCallback callback);

  using ProductsOfPoint = pdlp_b200::ProductsOfPoint;
  using ProductsOfCheck = pdlp_b200::ProductsOfCheck;
  // pdhg.cc:1434-1518.  Points are device pointers; avg / delta may be null.
  std::optional<TerminationReasonAndPointType> UpdateIterationStatsAndCheckTermination(
      const PrimalDualHybridGradientParams& params, bool force_numerical_termination, const double* x_cur, const double* y_cur,
      const double* x_avg, const double* y_avg, const double* dx, const double* dy, const double* x_start, const double* y_start,
      const InterruptFlag* interrupt, IterationType type, const IterationStats& full_stats, IterationStats& stats,
      const ProductsOfCheck* products = nullptr);
  // pdhg.cc:1520-1564.
  void ComputeConvergenceAndInfeasibility(const PrimalDualHybridGradientParams& params, const double* x, const double* y, const double* x_start,
                                          const double* y_start, int candidate_type, ConvergenceInformation* ci, InfeasibilityInformation* ii,
                                          PointMetadata* md, ProductsOfPoint products = {});
  // pdhg.cc:1568-1643.  x, y are device pointers into the scaled space.
  SolverResult ConstructOriginalSolverResult(const PrimalDualHybridGradientParams& params, const double* x, const double* y, SolveLog log);

  QuadraticProgramStats ComputeStats();
  // Sharded solves: the largest `code` over all ranks (every rank must call it).
  int AgreeOnCode(int code);
  // Max over ranks of each of `count` (<= 8) non-negative values, in place.
  void AgreeOnMax(double* values, int count);
  bool sharded() const { return comm_.sharded(); }
  // Kernel-level entry (pdlp_b200_rescale): ApplyRescaling on the resident problem.
  void RescaleForTest(const PrimalDualHybridGradientParams& params) { ApplyRescaling(params); }
  Workspace& ws() { return ws_; }
  DeviceProblem& problem() { return problem_; }
  const QuadraticProgramBoundNorms& OriginalBoundNorms() const { return original_bound_norms_; }
  const std::vector<int32_t>& RandomProjectionSeeds() const { return seeds_; }

 private:
  void ApplyRescaling(const PrimalDualHybridGradientParams& params);
  void AddRandomProjections(const double* x, const double* y, PointMetadata* md);

  QpView view_;
  Comm comm_;
  Workspace ws_;
  DeviceProblem problem_;
  QuadraticProgramBoundNorms original_bound_norms_;
  IterationCallback callback_;
  std::vector<int32_t> seeds_;
  SolutionSink* sink_ = nullptr;
  PresolveHook* presolve_ = nullptr;
  // Unscales (x, y) of this problem, runs the postsolve hook and leaves the projected original-space point in
  // presolve_->x / y (RecoverOriginalSolution, pdhg.cc:1288-1341).
  void RecoverOriginalOnDevice(const double* x, const double* y);
  // Text logging (pdlp_logging.h): a table header every 15 lines, lines at most every log_interval_seconds; rank 0 only.
  int log_counter_ = 0;
  double time_of_last_log_ = -1.0e300;
 public:
  bool Logs(const PrimalDualHybridGradientParams& params, int level) const { return params.verbosity_level >= level && comm_.rank() == 0; }
};

enum class InnerStepOutcome { kSuccessful, kForceNumericalTermination };

// pdhg.cc:504-709.  Owns the iterate buffers and the device-resident step state.
class Solver {
 public:
  // x0 / y0 are device pointers (scaled space), copied into the iterate buffers.
  Solver(const PrimalDualHybridGradientParams& params, const double* x0, const double* y0, double initial_step_size, double initial_primal_weight,
         PreprocessSolver* preprocess_solver);
  // Returns the log plus the device pointers of the chosen point (valid until the
  // Solver is destroyed).
  struct Outcome {
    SolveLog log;
    const double* x = nullptr;
    const double* y = nullptr;
  };
  Outcome Solve(IterationType type, const InterruptFlag* interrupt, SolveLog log);
  ~Solver();
  Solver(const Solver&) = delete;
  Solver& operator=(const Solver&) = delete;
  long long TrialSteps() const { return host_.total_trials; }
  // Times the device-resident loop stopped because the decision kernel saw the interrupt flag.
  int InterruptsSeenOnDevice() const { return interrupts_seen_on_device_; }
  double DeviceLoopSeconds() const { return device_loop_seconds_; }

 private:
  struct Bounds { double lagrangian_value, lower_bound, upper_bound, radius; };
  const double* X() const { return it_.x[host_.cur].get(); }
  const double* Y() const { return it_.y[host_.cur].get(); }
  const double* Kty() const { return it_.kty[host_.cur].get(); }
  bool AverageHasWeight() const { return host_.avg_weight_sum > 0.0; }
  const double* PrimalAverage() const { return host_.avg_weight_sum > 0.0 ? it_.avg_x.get() : X(); }
  const double* DualAverage() const { return host_.avg_weight_sum_dual > 0.0 ? it_.avg_y.get() : Y(); }

  void PushState();
  void PullState();
  void RunDeviceIterations(int target_iteration);
  int NextEventIteration(int k, IterationType type, int next_polishing_iteration) const;

  IterationStats CreateSimpleIterationStats(int restart_used) const;
  double DistanceTraveledFromLastStart(const double* x, const double* y);
  Bounds BoundsAt(const double* x, const double* y, const double* kty_or_null, const double* kx_or_null);
  // Host events (one GPU): K x of the current iterate, the average and the last difference, K^T y of the average and
  // the difference - everything the restart bounds and the termination statistics of this event multiply by K - from
  // ONE pass over the CSR copy and ONE over the CSC copy (pdlp_kkt.cu:ComputeProducts).  K^T y of the current iterate
  // is the cached one.  Valid until the iterate changes (a restart, the next trial steps).
  struct EventProducts {
    bool valid = false;
    const double* kx_cur = nullptr;
    const double* kx_avg = nullptr; const double* kty_avg = nullptr;    // null: no average at this event
    const double* kx_diff = nullptr; const double* kty_diff = nullptr;  // null: not a termination check, or no accepted step yet
  } event_;
  DVec event_m_[3], event_n_[2];
  void PrepareEventProducts(const double* dx, const double* dy);
  Bounds BoundsAtCurrent();
  Bounds BoundsAtAverage();
  bool ShouldDoAdaptiveRestartHeuristic(double candidate_normalized_gap) const;
  int DetermineDistanceBasedRestartChoice();
  int ChooseRestartToApply(bool is_major_iteration);
  double ComputeNewPrimalWeight();
  void ApplyRestartChoice(int restart_to_apply);
  std::optional<Outcome> MajorIterationAndTerminationCheck(IterationType type, bool force_numerical_termination, const InterruptFlag* interrupt,
                                                            const IterationStats& work_from_feasibility_polishing, SolveLog& log);
  Outcome PickSolutionAndConstructOutcome(const IterationStats& stats, int reason, int point_type, SolveLog log);
  IterationStats TotalWorkSoFar(const SolveLog& log) const;
  std::optional<Outcome> TryFeasibilityPolishing(int iteration_limit, const InterruptFlag* interrupt, SolveLog& log);

  const PrimalDualHybridGradientParams params_;
  PreprocessSolver* pre_;
  Workspace& ws_;
  DeviceProblem& problem_;
  Iterate it_;
  StepState host_{};  // mirror of the device state, valid after PullState()
  int iterations_completed_ = 0;
  bool have_delta_ = false;
  int kty_of_average_at_ = -1;  // iteration at which BoundsAtAverage left K^T y_avg in ws_.scratch_n[1] (-1: not there)
  double normalized_gap_at_last_trial_ = std::numeric_limits<double>::infinity();
  double normalized_gap_at_last_restart_ = std::numeric_limits<double>::infinity();
  double preprocessing_time_sec_ = 0.0;
  WallTimer timer_;
  struct { double distance_moved_last_restart_period = std::numeric_limits<double>::infinity(); int length_of_last_restart_period = 1; } distance_info_;
  int poll_iterations_ = 8;
  double kkt_passes_from_polishing_ = 0.0;  // cumulative_kkt_matrix_passes of the polishing phases so far (counts against kkt_matrix_pass_limit)
  bool poll_always_ = false;
  // Sharded solves: host decisions must be identical on every rank, so the two rank-local
  // inputs (wall clock, interrupt flag) are agreed on (max over ranks) once per host event.
  void AgreeOnClockAndInterrupt(const InterruptFlag* interrupt);
  double Seconds() const { return pre_->sharded() ? agreed_seconds_ : preprocessing_time_sec_ + timer_.Get(); }
  double agreed_seconds_ = 0.0;
  int agreed_seconds_iteration_ = 0;
  InterruptFlag agreed_interrupt_;
  double device_loop_seconds_ = 0.0;
  // Per-iteration interrupt (termination.cc:178; one GPU): the caller's flag lives in memory the device cannot be
  // assumed to read, so the solver owns a host-mapped pinned int that the decision kernel reads after every accepted
  // step; a watcher thread copies the caller's flag into it for the duration of Solve().
  void EnableDeviceInterrupt(const InterruptFlag* interrupt);
  void StopWatcher();
  const InterruptFlag* interrupt_ = nullptr;
  int* mapped_interrupt_ = nullptr;
  const volatile int* outer_stop_enqueue_ = nullptr;
  cudaEvent_t chunk_done_[2] = {nullptr, nullptr};  // throttle of the blind enqueue while an interrupt flag is live
  std::thread watcher_;
  std::atomic<bool> watcher_stop_{false};
  int interrupts_seen_on_device_ = 0;
  // PDLP_B200_DEBUG_TIMING: host wall time of the loop by phase (every phase ends in a device sync), printed at the end of Solve().
  struct { double steps = 0, restart_choice = 0, termination = 0, apply_restart = 0; int events = 0, checks = 0, restarts = 0, bounds = 0, tr_passes = 0; } phase_;
  // Holders for the result of feasibility polishing (device vectors must outlive Solve()).
  DVec polished_x_, polished_y_;
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_DRIVER_H_
