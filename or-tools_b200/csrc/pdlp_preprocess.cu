// PreprocessSolver: upload, bound clean-up, statistics, input screening, rescaling, the
// initial step size / primal weight, termination checks and result construction.
//   follows ortools/pdlp/primal_dual_hybrid_gradient.cc:711-1105, 1204-1286, 1414-1643.
// Glop presolve (:1107-1202) is host-only and not part of this path.
#include <exception>
#include <random>

#include <chrono>

#include "pdlp_driver.h"
#include "pdlp_logging.h"
#include "pdlp_launch.cuh"

namespace pdlp_b200 {

namespace {

std::string SixDigits(double v) {
  char b[64];
  std::snprintf(b, sizeof b, "%g", v);
  return b;
}

// pdhg.cc:731-897 (the log warnings are dropped; every rejection is kept).
std::optional<SolverResult> CheckProblemStats(const QuadraticProgramStats& s, bool check_excessively_small_values) {
  const double kBig = 1e50, kSmall = 1e-50;
  auto err = [](const std::string& m) { return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, m); };
  if (std::isnan(s.constraint_matrix_l2_norm.get())) return err("Constraint matrix has a NAN.");
  if (s.constraint_matrix_abs_max > kBig)
    return err("Constraint matrix has a non-zero with absolute value " + SixDigits(s.constraint_matrix_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (s.constraint_matrix_col_min_l_inf_norm > 0 && s.constraint_matrix_col_min_l_inf_norm < kSmall)
    return err("Constraint matrix has a column with Linf norm " + SixDigits(s.constraint_matrix_col_min_l_inf_norm) + " which is less than limit of " + SixDigits(kSmall) + ".");
  if (s.constraint_matrix_row_min_l_inf_norm > 0 && s.constraint_matrix_row_min_l_inf_norm < kSmall)
    return err("Constraint matrix has a row with Linf norm " + SixDigits(s.constraint_matrix_row_min_l_inf_norm) + " which is less than limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.combined_bounds_l2_norm.get())) return err("Constraint bounds vector has a NAN.");
  if (s.combined_bounds_max > kBig)
    return err("Combined constraint bounds vector has a non-zero with absolute value " + SixDigits(s.combined_bounds_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_excessively_small_values && s.combined_bounds_min > 0 && s.combined_bounds_min < kSmall)
    return err("Combined constraint bounds vector has a non-zero with absolute value " + SixDigits(s.combined_bounds_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.combined_variable_bounds_l2_norm.get())) return err("Variable bounds vector has a NAN.");
  if (s.combined_variable_bounds_max > kBig)
    return err("Combined variable bounds vector has a non-zero with absolute value " + SixDigits(s.combined_variable_bounds_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_excessively_small_values && s.combined_variable_bounds_min > 0 && s.combined_variable_bounds_min < kSmall)
    return err("Combined variable bounds vector has a non-zero with absolute value " + SixDigits(s.combined_variable_bounds_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.objective_vector_l2_norm.get())) return err("Objective vector has a NAN.");
  if (s.objective_vector_abs_max > kBig)
    return err("Objective vector has a non-zero with absolute value " + SixDigits(s.objective_vector_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  if (check_excessively_small_values && s.objective_vector_abs_min > 0 && s.objective_vector_abs_min < kSmall)
    return err("Objective vector has a non-zero with absolute value " + SixDigits(s.objective_vector_abs_min) + " which is less than the limit of " + SixDigits(kSmall) + ".");
  if (std::isnan(s.objective_matrix_l2_norm.get())) return err("Objective matrix has a NAN.");
  if (s.objective_matrix_abs_max > kBig)
    return err("Objective matrix has a non-zero with absolute value " + SixDigits(s.objective_matrix_abs_max) + " which exceeds limit of " + SixDigits(kBig) + ".");
  return std::nullopt;
}

}  // namespace

SolverResult ErrorSolverResult(int reason, const std::string& message) {
  SolverResult r;
  r.solve_log.termination_reason = reason;
  r.solve_log.termination_string = message;
  return r;
}

PreprocessSolver::PreprocessSolver(const QpView& qp, const PrimalDualHybridGradientParams& params) : view_(qp) {
  const TraceRange trace("pdlp/setup/upload_and_transpose");
  // num_threads / num_shards (pdhg.cc:71-96) configure the reference's CPU thread pool;
  // the device path accepts and ignores them.
  ws_.Init(qp.n, qp.m);
  if (qp.world > 1) {
    if (qp.n < 1 || qp.m < 1) throw std::invalid_argument("pdlp_b200: a sharded solve needs at least one column per rank and one constraint");
    comm_.Init(qp.rank, qp.world, qp.nccl_unique_id);
    ws_.comm = &comm_;
    ws_.shared_rows = qp.num_shared_rows < 0 ? qp.m : qp.num_shared_rows;
    ws_.dual_skip_below = qp.rank == 0 ? 0 : ws_.shared_rows;
    // Exchange buffers have the same size on every rank (NCCL symmetric registration needs that).
    double most_rows = static_cast<double>(qp.m);
    AgreeOnMax(&most_rows, 1);
    ws_.exchange_rows = comm_.ExchangeBufferFor(0, static_cast<size_t>(most_rows) * sizeof(double));
    ws_.exchange_sums = comm_.ExchangeBufferFor(1, 16 * sizeof(double));
    // The shared constraints are the same ones on every rank (include/pdlp_b200.h); the peer-memory slots
    // are sized by their number.  A disagreement would corrupt the exchange, so it is checked.
    double agreed[2] = {static_cast<double>(ws_.shared_rows), -static_cast<double>(ws_.shared_rows)};
    AgreeOnMax(agreed, 2);
    if (agreed[0] != -agreed[1]) throw std::invalid_argument("pdlp_b200: num_shared_constraints differs between ranks");
    ws_.peer = comm_.SetupPeerExchange(ws_.shared_rows, ws_.stream);
  }
  problem_.n = qp.n;
  problem_.m = qp.m;
  problem_.is_lp = qp.q == nullptr;
  problem_.objective_offset = qp.objective_offset;
  problem_.objective_scaling_factor = qp.objective_scaling_factor;
  // Sharded: the CSR copy's work blocks are cut at the shared / private row boundary (pdlp_launch.cuh:EnqueueRows).
  const int64_t row_cut = comm_.sharded() ? ws_.shared_rows : -1;
  // A failure on one rank only (a malformed block, out of memory) must not leave the other ranks waiting in the next
  // collective: the ranks agree on whether every block was built before anyone goes on.
  std::exception_ptr build_error;
  try {
    if (qp.col_ptr32 != nullptr) BuildDeviceMatrix(qp.n, qp.m, qp.nnz, qp.col_ptr32, qp.row_idx32, qp.val, ws_.stream, &problem_.K, row_cut);
    else BuildDeviceMatrix(qp.n, qp.m, qp.nnz, qp.col_ptr, qp.row_idx, qp.val, ws_.stream, &problem_.K, row_cut);
  } catch (...) {
    if (!comm_.sharded()) throw;
    build_error = std::current_exception();
  }
  if (comm_.sharded() && AgreeOnCode(build_error ? 1 : 0) != 0) {
    if (build_error) std::rethrow_exception(build_error);
    throw std::runtime_error("pdlp_b200: another rank could not build its block of the constraint matrix");
  }
  problem_.K.by_col.sharded_rows = comm_.sharded();
  problem_.K.by_row.partial_sums = comm_.sharded();
  auto up = [&](DVec& d, const double* h, int64_t count) {
    d.Resize(count > 0 ? count : 1);
    d.Upload(h, count, ws_.stream);
  };
  up(problem_.c, qp.c, qp.n);
  if (!problem_.is_lp) up(problem_.q, qp.q, qp.n);
  up(problem_.vlb, qp.vlb, qp.n);
  up(problem_.vub, qp.vub, qp.n);
  up(problem_.clb, qp.clb, qp.m);
  up(problem_.cub, qp.cub, qp.m);
  problem_.col_scale.Resize(qp.n > 0 ? qp.n : 1);
  problem_.row_scale.Resize(qp.m > 0 ? qp.m : 1);
  FillVector(ws_, problem_.col_scale.get(), qp.n, 1.0);
  FillVector(ws_, problem_.row_scale.get(), qp.m, 1.0);
  seeds_ = params.random_projection_seeds.v;
  ws_.Sync();
}

PreprocessSolver::~PreprocessSolver() {
  if (ws_.stream != nullptr) cudaStreamSynchronize(ws_.stream);
  comm_.Destroy();
  ws_.Destroy();
}

void PreprocessSolver::AgreeOnMax(double* values, int count) {
  if (!comm_.sharded()) return;
  PDLP_CUDA_CHECK(cudaMemcpyAsync(ws_.totals.get(), values, count * sizeof(double), cudaMemcpyHostToDevice, ws_.stream));
  comm_.AllReduce(ws_.totals.get(), count, true, ws_.stream);
  const double* agreed = FetchTotals(ws_, count);
  for (int i = 0; i < count; ++i) values[i] = agreed[i];
}
int PreprocessSolver::AgreeOnCode(int code) {
  double v = code;
  AgreeOnMax(&v, 1);
  return static_cast<int>(v);
}

// sharded_optimization_utils.cc:270-343.
QuadraticProgramStats PreprocessSolver::ComputeStats() {
  const TraceRange trace("pdlp/setup/stats");
  DeviceProblem& p = problem_;
  QuadraticProgramStats s;
  // Row / column Linf norms with unit scaling (ones live in scratch).
  double* ones_n = ws_.scratch_n[1].get();
  double* ones_m = ws_.scratch_m[1].get();
  FillVector(ws_, ones_n, p.n, 1.0);
  FillVector(ws_, ones_m, p.m, 1.0);
  double* row_norms = ws_.scratch_m[2].get();
  double* col_norms = ws_.scratch_n[2].get();
  ScaledRowNorms(ws_, p.K.by_row, ones_n, ones_m, false, row_norms);
  ScaledRowNorms(ws_, p.K.by_col, ones_m, ones_n, false, col_norms);
  const VectorInfo rn = ComputeVectorInfo(ws_, InfoKind::kPlain, row_norms, nullptr, p.m, Side::kDual);
  const VectorInfo cn = ComputeVectorInfo(ws_, InfoKind::kPlain, col_norms, nullptr, p.n, Side::kSharded);
  const VectorInfo mat = ComputeVectorInfo(ws_, InfoKind::kPlain, p.K.by_col.val.get(), nullptr, p.K.by_col.nnz, Side::kSharded);
  const VectorInfo cb = ComputeVectorInfo(ws_, InfoKind::kCombinedBounds, p.cub.get(), p.clb.get(), p.m, Side::kDual);
  const VectorInfo vb = ComputeVectorInfo(ws_, InfoKind::kCombinedBounds, p.vub.get(), p.vlb.get(), p.n, Side::kSharded);
  const VectorInfo ov = ComputeVectorInfo(ws_, InfoKind::kPlain, p.c.get(), nullptr, p.n, Side::kSharded);
  const VectorInfo gaps = ComputeVectorInfo(ws_, InfoKind::kGap, p.vlb.get(), p.vub.get(), p.n, Side::kSharded);
  s.num_variables = comm_.sharded() ? ov.num_finite_nonzero + ov.num_zero + ov.num_infinite : p.n;  // global count
  s.num_constraints = comm_.sharded() ? cb.num_finite_nonzero + cb.num_zero + cb.num_infinite : p.m;  // global count
  s.constraint_matrix_col_min_l_inf_norm = cn.smallest;
  s.constraint_matrix_row_min_l_inf_norm = rn.smallest;
  s.constraint_matrix_num_nonzeros = mat.num_finite_nonzero;
  s.constraint_matrix_abs_max = mat.largest;
  s.constraint_matrix_abs_min = mat.smallest;
  s.constraint_matrix_abs_avg = mat.average;
  s.constraint_matrix_l2_norm = mat.l2_norm;
  s.combined_bounds_max = cb.largest;
  s.combined_bounds_min = cb.smallest;
  s.combined_bounds_avg = cb.average;
  s.combined_bounds_l2_norm = cb.l2_norm;
  s.combined_variable_bounds_max = vb.largest;
  s.combined_variable_bounds_min = vb.smallest;
  s.combined_variable_bounds_avg = vb.average;
  s.combined_variable_bounds_l2_norm = vb.l2_norm;
  s.variable_bound_gaps_num_finite = gaps.num_finite_nonzero + gaps.num_zero;
  s.variable_bound_gaps_max = gaps.largest;
  s.variable_bound_gaps_min = gaps.smallest;
  s.variable_bound_gaps_avg = gaps.average;
  s.variable_bound_gaps_l2_norm = gaps.l2_norm;
  s.objective_vector_abs_max = ov.largest;
  s.objective_vector_abs_min = ov.smallest;
  s.objective_vector_abs_avg = ov.average;
  s.objective_vector_l2_norm = ov.l2_norm;
  if (p.is_lp) {
    s.objective_matrix_num_nonzeros = 0;
    s.objective_matrix_abs_max = 0;
    s.objective_matrix_abs_min = 0;
    s.objective_matrix_abs_avg = std::numeric_limits<double>::quiet_NaN();
    s.objective_matrix_l2_norm = 0;
  } else {
    const VectorInfo om = ComputeVectorInfo(ws_, InfoKind::kPlain, p.q.get(), nullptr, p.n, Side::kSharded);
    s.objective_matrix_num_nonzeros = om.num_finite_nonzero;
    s.objective_matrix_abs_max = om.largest;
    s.objective_matrix_abs_min = om.smallest;
    s.objective_matrix_abs_avg = om.average;
    s.objective_matrix_l2_norm = om.l2_norm;
  }
  return s;
}

// sharded_optimization_utils.cc:367-444 + sharded_quadratic_program.cc:134-167.
void PreprocessSolver::ApplyRescaling(const PrimalDualHybridGradientParams& params) {
  const TraceRange trace("pdlp/setup/rescaling");
  DeviceProblem& p = problem_;
  double* row = p.row_scale.get();
  double* col = p.col_scale.get();
  double* col_norm = ws_.scratch_n[2].get();
  double* row_norm = ws_.scratch_m[2].get();
  auto iteration = [&](bool l2) {
    ScaledRowNorms(ws_, p.K.by_col, row, col, l2, col_norm);
    ScaledRowNorms(ws_, p.K.by_row, col, row, l2, row_norm);
    DivideBySqrtOfDivisor(ws_, col_norm, col, p.n);
    DivideBySqrtOfDivisor(ws_, row_norm, row, p.m);
  };
  bool any = false;
  for (int i = 0; i < params.l_inf_ruiz_iterations; ++i) { iteration(false); any = true; }
  if (params.l2_norm_rescaling) { iteration(true); any = true; }
  if (any) {
    ScaleProblemVectors(ws_, p, col, row);
    ScaleMatrixValues(ws_, p.K.by_col, row, col);
    ScaleMatrixValues(ws_, p.K.by_row, col, row);
  }
}

// pdhg.cc:949-1105.
SolverResult PreprocessSolver::PreprocessAndSolve(const PrimalDualHybridGradientParams& params, const double* x0, const double* y0,
                                                  const InterruptFlag* interrupt, IterationCallback callback) {
  WallTimer timer;
  timer.Start();
  SolveLog log;
  if (view_.name != nullptr) log.instance_name = std::string(view_.name);
  log.params.mut() = params;
  DeviceProblem& p = problem_;
  ReplaceLargeBoundsWithInfinity(ws_, p.clb.get(), p.cub.get(), p.m, params.infinite_constraint_bound_threshold);
  if (presolve_ != nullptr) {
    // The statistics of the ORIGINAL problem (pdhg.cc:962-978 run before ApplyPresolveIfEnabled); a warm start
    // cannot be mapped into the presolved space and is dropped with the reference's warning (:1164-1168).
    PreprocessSolver& o = *presolve_->original;
    ReplaceLargeBoundsWithInfinity(o.ws_, o.problem_.clb.get(), o.problem_.cub.get(), o.problem_.m, params.infinite_constraint_bound_threshold);
    log.original_problem_stats.mut() = o.ComputeStats();
    if (x0 != nullptr || y0 != nullptr) {
      std::fputs("WARNING: Ignoring initial solution. Initial solutions are ignored when presolve is on.\n", stderr);
      x0 = y0 = nullptr;
    }
  } else {
    log.original_problem_stats.mut() = ComputeStats();
  }
  if (auto r = CheckProblemStats(log.original_problem_stats.get(), presolve_ != nullptr || params.presolve_options->use_glop); r.has_value()) return *r;

  // Starting point (pdhg.cc:1035-1048).  The size checks of CheckInitialSolution
  // (:899-947) happen at the API boundary where the sizes are known.
  DVec x_start(p.n > 0 ? p.n : 1), y_start(p.m > 0 ? p.m : 1);
  if (x0 != nullptr && y0 != nullptr) {
    x_start.Upload(x0, p.n, ws_.stream);
    y_start.Upload(y0, p.m, ws_.stream);
    const double kBig = 1e50;
    double l2 = 0, linf = 0;
    NormAndMax(ws_, x_start.get(), p.n, &l2, &linf, Side::kSharded);
    if (std::isnan(l2)) return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION, "Initial primal solution has a NAN.");
    if (linf > kBig)
      return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION,
                               "Initial primal solution has an entry with absolute value " + SixDigits(linf) + " which exceeds limit of " + SixDigits(kBig));
    NormAndMax(ws_, y_start.get(), p.m, &l2, &linf, Side::kDual);
    if (std::isnan(l2)) return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION, "Initial dual solution has a NAN.");
    if (linf > kBig)
      return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION,
                               "Initial dual solution has an entry with absolute value " + SixDigits(linf) + " which exceeds limit of " + SixDigits(kBig));
  } else {
    x_start.Zero(ws_.stream);
    y_start.Zero(ws_.stream);
  }
  original_bound_norms_ = BoundNormsFromProblemStats(log.original_problem_stats.get());
  if (Logs(params, 1)) {  // pdhg.cc:979-986
    LogInfoWithoutPrefix("Problem stats before rescaling:");
    for (const std::string& line : QuadraticProgramStatsLines(log.original_problem_stats.get(), log.original_problem_stats->constraint_matrix_num_nonzeros > 0, problem_.is_lp)) LogInfoWithoutPrefix(line);
  }
  callback_ = std::move(callback);

  ProjectToPrimalBounds(ws_, p, x_start.get());
  ProjectToDualBounds(ws_, p, y_start.get());

  ApplyRescaling(params);  // pdhg.cc:1204-1218
  DivideInPlace(ws_, x_start.get(), p.col_scale.get(), p.n);
  DivideInPlace(ws_, y_start.get(), p.row_scale.get(), p.m);
  log.preprocessed_problem_stats.mut() = ComputeStats();
  if (Logs(params, 1)) {  // pdhg.cc:1053-1057
    LogInfoWithoutPrefix("Problem stats after rescaling:");
    for (const std::string& line : QuadraticProgramStatsLines(log.preprocessed_problem_stats.get(), log.preprocessed_problem_stats->constraint_matrix_num_nonzeros > 0, problem_.is_lp)) LogInfoWithoutPrefix(line);
  }

  double step_size = 0.0;
  if (params.linesearch_rule == CONSTANT_STEP_SIZE_RULE) {
    // pdhg.cc:1060-1076 + sou.cc:536-699 (power method on K^T K, 2 SpMV per round).
    // parity unpinned for the exact estimate: the reference draws its start vector with
    // absl::Gaussian(mt19937(1)), whose algorithm is not in the reference tree.
    const double desired_relative_error = 0.2, failure_probability = 0.0005;
    std::mt19937 random(1);
    std::normal_distribution<double> gauss(0.0, 1.0);
    // One global sequence; a sharded rank keeps its own slice of it.
    std::vector<double> start(p.n);
    for (int64_t j = 0; j < view_.col_offset; ++j) (void)gauss(random);
    for (double& v : start) v = gauss(random);
    double* ev = ws_.scratch_n[0].get();
    double* next = ws_.scratch_n[1].get();
    double* dual_ev = ws_.scratch_m[0].get();
    ws_.scratch_n[0].Upload(start.data(), p.n, ws_.stream);
    auto normalize = [&](double* v) {
      double l2 = 0, linf = 0;
      NormAndMax(ws_, v, p.n, &l2, &linf, Side::kSharded);
      if (l2 != 0.0) {
        FillVector(ws_, ws_.scratch_n[2].get(), p.n, l2);
        DivideInPlace(ws_, v, ws_.scratch_n[2].get(), p.n);
      }
    };
    normalize(ev);
    const double eps = 1.0 - (1.0 - desired_relative_error) * (1.0 - desired_relative_error);
    auto fail_prob = [&](int k) {
      if (k < 2 || eps <= 0.0) return 1.0;
      return std::min(0.824, 0.354 / std::sqrt(eps * (k - 1))) * std::sqrt(static_cast<double>(view_.GlobalN())) * std::pow(1.0 - eps, k - 0.5);
    };
    double estimate = 0.0;
    int iters = 0;
    while (fail_prob(iters) > failure_probability) {
      RowDot(ws_, p.K.by_row, ev, dual_ev);
      RowDot(ws_, p.K.by_col, dual_ev, next);
      // estimate = ev . next = |K ev|^2
      double l2 = 0, linf = 0;
      NormAndMax(ws_, dual_ev, p.m, &l2, &linf, Side::kDual);
      estimate = l2 * l2;
      CopyVector(ws_, ev, next, p.n);
      ++iters;
      normalize(ev);
    }
    const double singular_value = std::sqrt(estimate);
    const double inverse_step_lower_bound = singular_value / (1.0 - desired_relative_error);
    step_size = inverse_step_lower_bound > 0.0 ? 1.0 / inverse_step_lower_bound : 1.0;
  } else {
    step_size = 1.0 / std::max(1.0e-20, log.preprocessed_problem_stats->constraint_matrix_abs_max.get());
  }
  step_size *= params.initial_step_size_scaling;

  // InitialPrimalWeight, pdhg.cc:1268-1286.
  double primal_weight = 1.0;
  const double c_norm = log.preprocessed_problem_stats->objective_vector_l2_norm;
  const double b_norm = log.preprocessed_problem_stats->combined_bounds_l2_norm;
  if (params.initial_primal_weight.has) primal_weight = params.initial_primal_weight;
  else if (c_norm > 0.0 && b_norm > 0.0) primal_weight = c_norm / b_norm;

  Solver solver(params, x_start.get(), y_start.get(), step_size, primal_weight, this);
  log.preprocessing_time_sec = timer.Get();
  Solver::Outcome outcome = solver.Solve(IterationType::kNormal, interrupt, std::move(log));
  SolverResult result = ConstructOriginalSolverResult(params, outcome.x, outcome.y, std::move(outcome.log));
  // Counters for benchmarks.
  g_last_counters.kernel_launches = ws_.launches;
  g_last_counters.collectives = comm_.collectives() + comm_.peer_exchanges();  // exchange steps: NCCL all-reduces + peer-memory exchanges
  g_last_counters.symmetric_buffers = comm_.symmetric_buffers();
  g_last_counters.peer_exchanges = comm_.peer_exchanges();
  g_last_counters.trial_steps = solver.TrialSteps();
  g_last_counters.interrupts_seen_on_device = solver.InterruptsSeenOnDevice();
  g_last_counters.device_loop_seconds = solver.DeviceLoopSeconds();
  g_last_counters.setup_seconds = result.solve_log.preprocessing_time_sec;
  g_last_counters.loop_seconds = result.solve_log.solve_time_sec - result.solve_log.preprocessing_time_sec;
  return result;
}

// pdhg.cc:1520-1564 + iteration_stats.cc:349-518, assembled from the fused totals.
void PreprocessSolver::ComputeConvergenceAndInfeasibility(const PrimalDualHybridGradientParams& params, const double* x, const double* y,
                                                          const double* x_start, const double* y_start, int candidate_type,
                                                          ConvergenceInformation* ci, InfeasibilityInformation* ii, PointMetadata* md,
                                                          ProductsOfPoint products) {
  const DetailedOptimalityCriteria oc = EffectiveOptimalityCriteria(params.termination_criteria.get());
  const double primal_ratio = EpsilonRatio(oc.eps_optimal_primal_residual_absolute, oc.eps_optimal_primal_residual_relative);
  const double dual_ratio = EpsilonRatio(oc.eps_optimal_dual_residual_absolute, oc.eps_optimal_dual_residual_relative);
  if (presolve_ != nullptr && (ci != nullptr || ii != nullptr)) {
    // pdhg.cc:1530-1548: under presolve the convergence and infeasibility information are those of the ORIGINAL
    // problem (trivial scaling) at the recovered point; the point metadata stay those of the working problem.
    RecoverOriginalOnDevice(x, y);
    presolve_->original->ComputeConvergenceAndInfeasibility(params, presolve_->x.get(), presolve_->y.get(), presolve_->x.get(), presolve_->y.get(), candidate_type, ci,
                                                            ii, nullptr);
    if (md == nullptr) return;
    ci = nullptr;
    ii = nullptr;
  }
  const bool flag = params.handle_some_primal_gradients_on_finite_bounds_as_residuals;
  const bool have_products = !comm_.sharded() && (products.kx != nullptr || problem_.m == 0) && (products.kty != nullptr || problem_.n == 0);
  const KktTotals t = have_products ? ComputeKktTotalsFromProducts(ws_, problem_, x, y, x_start, y_start, products.kx, products.kty, primal_ratio, dual_ratio, flag)
                                    : ComputeKktTotals(ws_, problem_, x, y, x_start, y_start, primal_ratio, dual_ratio, flag);
  auto apply = [&](double v) { return problem_.objective_scaling_factor * (v + problem_.objective_offset); };
  if (ci != nullptr) {
    ci->l_inf_primal_residual = t.primal_res_linf;
    ci->l2_primal_residual = std::sqrt(t.primal_res_sumsq);
    ci->l_inf_componentwise_primal_residual = t.primal_res_cw;
    ci->l_inf_primal_variable = t.x_linf;
    ci->l2_primal_variable = std::sqrt(t.x_sumsq);
    ci->l_inf_dual_variable = t.y_linf;
    ci->l2_dual_variable = std::sqrt(t.y_sumsq);
    const double quad = problem_.is_lp ? 0.0 : 0.5 * t.x_q_x;
    ci->primal_objective = apply(quad + t.c_dot_x);
    const double dual_piece = -quad + t.dual_bounds_term;
    ci->dual_objective = apply(dual_piece + t.correction);
    ci->corrected_dual_objective = apply(dual_piece + t.full_correction);
    ci->l_inf_dual_residual = t.dual_res_linf;
    ci->l2_dual_residual = std::sqrt(t.dual_res_sumsq);
    ci->l_inf_componentwise_dual_residual = t.dual_res_cw;
    ci->candidate_type = candidate_type;
  }
  if (ii != nullptr) {
    const double linf_primal = t.x_linf, linf_dual = t.y_linf;
    ii->max_primal_ray_infeasibility = 0.0; ii->primal_ray_linear_objective = 0.0; ii->primal_ray_quadratic_norm = 0.0;
    ii->max_dual_ray_infeasibility = 0.0; ii->dual_ray_objective = 0.0;
    const double dual_ray_objective = t.dual_bounds_term + t.ray_correction;
    if (linf_dual > 0) {
      ii->dual_ray_objective = dual_ray_objective / linf_dual;
      ii->max_dual_ray_infeasibility = t.ray_dual_res_linf / linf_dual;
    }
    if (linf_primal > 0.0) {
      ii->primal_ray_quadratic_norm = t.qx_linf / linf_primal;
      ii->max_primal_ray_infeasibility = std::max(t.primal_res_linf_homogeneous, t.sign_violation) / linf_primal;
      ii->primal_ray_linear_objective = t.c_dot_x / linf_primal;
    }
    ii->candidate_type = candidate_type;
  }
  if (md != nullptr) {
    md->point_type = candidate_type;
    if (candidate_type != POINT_TYPE_ITERATE_DIFFERENCE) {  // pdhg.cc:1343-1412
      md->active_primal_variable_count = static_cast<int64_t>(t.active_primal_count);
      md->active_dual_variable_count = static_cast<int64_t>(t.active_dual_count);
      md->active_primal_variable_change = static_cast<int64_t>(t.active_primal_change);
      md->active_dual_variable_change = static_cast<int64_t>(t.active_dual_change);
    }
    if (!seeds_.empty()) AddRandomProjections(x, y, md);  // host-side diagnostic (sharded: partial sums added over the ranks)
  }
}

// iteration_stats.cc:321-346,598-610.  Rarely used diagnostics: computed on the host
// from a copy of the point.  parity unpinned (absl::Gaussian; depends on the shard count
// by design, solvers.proto:397-403): one shard per GPU is used here.
// Column-sharded solves: the random direction is assembled from one Gaussian stream per owner - the shared rows'
// stream is the same on every rank (seeded without a rank) and counted on rank 0, every rank's columns and private
// rows draw from a stream seeded with its rank - and the partial dot products and squared norms are added over
// the ranks, so all ranks record the same projection of the WHOLE point onto one unit direction per seed.
void PreprocessSolver::AddRandomProjections(const double* x, const double* y, PointMetadata* md) {
  std::vector<double> hx(problem_.n), hy(problem_.m);
  if (problem_.n > 0) PDLP_CUDA_CHECK(cudaMemcpyAsync(hx.data(), x, problem_.n * sizeof(double), cudaMemcpyDeviceToHost, ws_.stream));
  if (problem_.m > 0) PDLP_CUDA_CHECK(cudaMemcpyAsync(hy.data(), y, problem_.m * sizeof(double), cudaMemcpyDeviceToHost, ws_.stream));
  ws_.Sync();
  // Adds the dot product of v[begin, end) with a Gaussian stream, and the stream's squared norm.
  auto accumulate = [](const std::vector<double>& v, int64_t begin, int64_t end, std::mt19937& rnd, double* dot, double* norm2) {
    std::normal_distribution<double> gauss(0.0, 1.0);
    for (int64_t i = begin; i < end; ++i) { const double g = gauss(rnd); *dot += g * v[i]; *norm2 += g * g; }
  };
  const bool sharded = comm_.sharded();
  const int64_t shared = sharded ? ws_.shared_rows : 0;
  std::vector<double> sums(4 * seeds_.size(), 0.0);  // per seed: primal dot, primal norm^2, dual dot, dual norm^2
  for (size_t k = 0; k < seeds_.size(); ++k) {
    std::mt19937 gen(seeds_[k]);
    const uint32_t primal_seed = gen(), dual_seed = gen();
    double* s = &sums[4 * k];
    if (!sharded) {
      std::mt19937 primal_rnd(primal_seed), dual_rnd(dual_seed);
      accumulate(hx, 0, problem_.n, primal_rnd, s + 0, s + 1);
      accumulate(hy, 0, problem_.m, dual_rnd, s + 2, s + 3);
      continue;
    }
    const uint32_t owner = static_cast<uint32_t>(comm_.rank()) + 1u;
    std::seed_seq primal_seq{primal_seed, owner}, private_seq{dual_seed, owner};
    std::mt19937 primal_rnd(primal_seq), private_rnd(private_seq), shared_rnd(dual_seed);
    accumulate(hx, 0, problem_.n, primal_rnd, s + 0, s + 1);
    if (comm_.rank() == 0) accumulate(hy, 0, shared, shared_rnd, s + 2, s + 3);
    accumulate(hy, shared, problem_.m, private_rnd, s + 2, s + 3);
  }
  if (sharded) {
    constexpr size_t kChunk = 32;  // doubles per all-reduce (ws_.totals holds kMaxSlots * 4)
    for (size_t at = 0; at < sums.size(); at += kChunk) {
      const size_t count = std::min(kChunk, sums.size() - at);
      PDLP_CUDA_CHECK(cudaMemcpyAsync(ws_.totals.get(), sums.data() + at, count * sizeof(double), cudaMemcpyHostToDevice, ws_.stream));
      comm_.AllReduce(ws_.totals.get(), count, false, ws_.stream);
      const double* total = FetchTotals(ws_, static_cast<int>(count));
      std::copy(total, total + count, sums.begin() + at);
    }
  }
  for (size_t k = 0; k < seeds_.size(); ++k) {
    const double* s = &sums[4 * k];
    md->random_primal_projections.v.push_back(s[0] / std::sqrt(s[1]));
    md->random_dual_projections.v.push_back(s[2] / std::sqrt(s[3]));
  }
}

// pdhg.cc:1434-1518.
std::optional<TerminationReasonAndPointType> PreprocessSolver::UpdateIterationStatsAndCheckTermination(
    const PrimalDualHybridGradientParams& params, bool force_numerical_termination, const double* x_cur, const double* y_cur, const double* x_avg,
    const double* y_avg, const double* dx, const double* dy, const double* x_start, const double* y_start, const InterruptFlag* interrupt,
    IterationType type, const IterationStats& full_stats, IterationStats& stats, const ProductsOfCheck* products) {
  auto add_point = [&](const double* x, const double* y, int point_type, bool convergence, ProductsOfPoint of_point) {
    ConvergenceInformation ci;
    InfeasibilityInformation ii;
    PointMetadata md;
    ComputeConvergenceAndInfeasibility(params, x, y, x_start, y_start, point_type, convergence ? &ci : nullptr, &ii, &md, of_point);
    if (convergence) stats.convergence_information.push_back(ci);
    stats.infeasibility_information.push_back(ii);
    stats.point_metadata.push_back(md);
  };
  const ProductsOfCheck none;
  const ProductsOfCheck& have = products != nullptr ? *products : none;
  add_point(x_cur, y_cur, POINT_TYPE_CURRENT_ITERATE, true, have.current);
  if (x_avg != nullptr && y_avg != nullptr) add_point(x_avg, y_avg, POINT_TYPE_AVERAGE_ITERATE, true, have.average);
  if (dx != nullptr && dy != nullptr) add_point(dx, dy, POINT_TYPE_ITERATE_DIFFERENCE, false, have.difference);
  if (Logs(params, 2)) {  // :1471-1501
    constexpr int kLogEvery = 15;
    const double now = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    if (params.log_interval_seconds == 0.0 || now - time_of_last_log_ >= params.log_interval_seconds) {
      const int level = params.verbosity_level;
      if (log_counter_ == 0) LogInfoWithoutPrefix(IterationStatsHeader(level, params.use_feasibility_polishing));
      LogInfoWithoutPrefix(IterationStatsLine(level, params.use_feasibility_polishing, type, stats, params.termination_criteria.get(), original_bound_norms_, POINT_TYPE_AVERAGE_ITERATE));
      // When there is no average the line above already shows the current iterate: it is not repeated.
      if (level >= 4 && GetConvergenceInformationFor(stats, POINT_TYPE_AVERAGE_ITERATE).has_value()) {
        LogInfoWithoutPrefix(IterationStatsLine(level, params.use_feasibility_polishing, type, stats, params.termination_criteria.get(), original_bound_norms_, POINT_TYPE_CURRENT_ITERATE));
      }
      time_of_last_log_ = now;
      if (++log_counter_ >= kLogEvery) log_counter_ = 0;
    }
  }
  if (callback_) callback_(IterationCallbackInfo{type, params.termination_criteria.get(), stats, original_bound_norms_});
  if (auto t = CheckIterateTerminationCriteria(params.termination_criteria.get(), stats, original_bound_norms_, force_numerical_termination); t.has_value()) return t;
  return CheckSimpleTerminationCriteria(params.termination_criteria.get(), full_stats, interrupt);
}

// RecoverOriginalSolution (pdhg.cc:1288-1341) with the caller's postsolve in the place of Glop's.
void PreprocessSolver::RecoverOriginalOnDevice(const double* x, const double* y) {
  DeviceProblem& p = problem_;
  PreprocessSolver& o = *presolve_->original;
  std::vector<double> hx(p.n), hy(p.m), ox(o.problem_.n), oy(o.problem_.m);
  if (p.n > 0) {
    double* t = ws_.scratch_n[2].get();
    CopyVector(ws_, t, x, p.n);
    MultiplyInPlace(ws_, t, p.col_scale.get(), p.n);
    PDLP_CUDA_CHECK(cudaMemcpyAsync(hx.data(), t, p.n * sizeof(double), cudaMemcpyDeviceToHost, ws_.stream));
  }
  if (p.m > 0) {
    double* t = ws_.scratch_m[2].get();
    CopyVector(ws_, t, y, p.m);
    MultiplyInPlace(ws_, t, p.row_scale.get(), p.m);
    PDLP_CUDA_CHECK(cudaMemcpyAsync(hy.data(), t, p.m * sizeof(double), cudaMemcpyDeviceToHost, ws_.stream));
  }
  ws_.Sync();
  presolve_->postsolve(hx.data(), hy.data(), ox.data(), oy.data());
  if (presolve_->x.size() == 0) {
    presolve_->x.Resize(o.problem_.n > 0 ? o.problem_.n : 1);
    presolve_->y.Resize(o.problem_.m > 0 ? o.problem_.m : 1);
  }
  presolve_->x.Upload(ox.data(), o.problem_.n, o.ws_.stream);
  presolve_->y.Upload(oy.data(), o.problem_.m, o.ws_.stream);
  // "Glop's preprocessor sometimes violates the primal bounds constraints. To be safe we project both." (:1332-1337)
  ProjectToPrimalBounds(o.ws_, o.problem_, presolve_->x.get());
  ProjectToDualBounds(o.ws_, o.problem_, presolve_->y.get());
  o.ws_.Sync();
}

// The presolver solved the problem completely (pdhg.cc:986-1034).
SolverResult PreprocessSolver::FinishPresolvedOnly(const PrimalDualHybridGradientParams& params, int presolve_reason, const InterruptFlag* interrupt,
                                                   IterationCallback callback) {
  WallTimer timer;
  timer.Start();
  SolveLog log;
  if (view_.name != nullptr) log.instance_name = std::string(view_.name);
  log.params.mut() = params;
  PreprocessSolver& o = *presolve_->original;
  ReplaceLargeBoundsWithInfinity(o.ws_, o.problem_.clb.get(), o.problem_.cub.get(), o.problem_.m, params.infinite_constraint_bound_threshold);
  log.original_problem_stats.mut() = o.ComputeStats();
  if (auto r = CheckProblemStats(log.original_problem_stats.get(), true); r.has_value()) return *r;
  original_bound_norms_ = BoundNormsFromProblemStats(log.original_problem_stats.get());
  callback_ = std::move(callback);
  log.preprocessed_problem_stats.mut() = ComputeStats();  // the (empty or unused) presolved problem; trivial scaling
  IterationStats stats;
  stats.cumulative_time_sec = timer.Get();
  log.preprocessing_time_sec = stats.cumulative_time_sec.get();
  // Zero working vectors (:992-995); their recovered image is the presolver's solution.
  DVec zx(problem_.n > 0 ? problem_.n : 1), zy(problem_.m > 0 ? problem_.m : 1);
  zx.Zero(ws_.stream);
  zy.Zero(ws_.stream);
  ConvergenceInformation ci;
  InfeasibilityInformation ii;
  ComputeConvergenceAndInfeasibility(params, zx.get(), zy.get(), zx.get(), zy.get(), POINT_TYPE_PRESOLVER_SOLUTION, &ci, &ii, nullptr);
  stats.convergence_information.push_back(ci);
  stats.infeasibility_information.push_back(ii);
  auto earned = CheckIterateTerminationCriteria(params.termination_criteria.get(), stats, original_bound_norms_, false);
  if (!earned.has_value()) earned = CheckSimpleTerminationCriteria(params.termination_criteria.get(), stats, interrupt);
  int reason;
  if (earned.has_value() && (earned->reason == TERMINATION_REASON_OPTIMAL || earned->reason == TERMINATION_REASON_PRIMAL_INFEASIBLE ||
                             earned->reason == TERMINATION_REASON_DUAL_INFEASIBLE)) {
    reason = earned->reason;
  } else if (presolve_reason == TERMINATION_REASON_OPTIMAL) {
    reason = TERMINATION_REASON_NUMERICAL_ERROR;
    std::fputs("WARNING: Presolve claimed to solve the LP optimally but the solution doesn't satisfy the optimality criteria.\n", stderr);
  } else {
    reason = presolve_reason;
  }
  log.iteration_count = stats.iteration_number;
  log.termination_reason = reason;
  log.solution_type = POINT_TYPE_PRESOLVER_SOLUTION;
  log.solve_time_sec = stats.cumulative_time_sec.get();
  log.solution_stats.mut() = stats;
  return ConstructOriginalSolverResult(params, zx.get(), zy.get(), std::move(log));
}

// pdhg.cc:1568-1643.
SolverResult PreprocessSolver::ConstructOriginalSolverResult(const PrimalDualHybridGradientParams& params, const double* x, const double* y, SolveLog log) {
  if (presolve_ != nullptr) {
    // :1571-1585: recover the original solution; the reduced costs are computed with respect to the original
    // problem.  The original's PreprocessSolver does both (its scaling vectors are all ones).
    RecoverOriginalOnDevice(x, y);
    PreprocessSolver& o = *presolve_->original;
    o.set_solution_sink(sink_);
    o.callback_ = callback_;
    o.original_bound_norms_ = original_bound_norms_;
    return o.ConstructOriginalSolverResult(params, presolve_->x.get(), presolve_->y.get(), std::move(log));
  }
  DeviceProblem& p = problem_;
  SolverResult result;
  const bool zero_objective = log.termination_reason == TERMINATION_REASON_PRIMAL_INFEASIBLE;
  double* out_x = ws_.scratch_n[0].get();
  double* out_rc = ws_.scratch_n[1].get();
  double* out_y = ws_.scratch_m[0].get();
  ComputeReducedCosts(ws_, p, x, y, zero_objective, params.handle_some_primal_gradients_on_finite_bounds_as_residuals, out_rc);
  CopyVector(ws_, out_x, x, p.n);
  CopyVector(ws_, out_y, y, p.m);
  MultiplyInPlace(ws_, out_x, p.col_scale.get(), p.n);
  MultiplyInPlace(ws_, out_y, p.row_scale.get(), p.m);
  DivideInPlace(ws_, out_rc, p.col_scale.get(), p.n);
  // Straight into the caller's arrays when there is a sink (C ABI), else into the result's vectors.
  auto download = [&](const double* device, int64_t count, std::vector<double>* own, double* sink, bool use_sink) {
    double* host = sink;
    if (!use_sink) {
      own->resize(count);
      host = own->data();
    }
    if (count == 0 || host == nullptr) return;
    const size_t bytes = static_cast<size_t>(count) * sizeof(double);
    if (bytes >= HostStager::kMinBytes && HostStager::Get().Download(host, device, bytes, ws_.stream)) return;
    PDLP_CUDA_CHECK(cudaMemcpyAsync(host, device, bytes, cudaMemcpyDeviceToHost, ws_.stream));
  };
  const bool use_sink = sink_ != nullptr;
  download(out_x, p.n, &result.primal_solution, use_sink ? sink_->primal : nullptr, use_sink);
  download(out_rc, p.n, &result.reduced_costs, use_sink ? sink_->reduced_costs : nullptr, use_sink);
  download(out_y, p.m, &result.dual_solution, use_sink ? sink_->dual : nullptr, use_sink);
  ws_.Sync();
  if (use_sink) sink_->filled = true;
  IterationType type = IterationType::kNormalTermination;
  if (log.solution_type == POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION) type = IterationType::kFeasibilityPolishingTermination;
  else if (log.solution_type == POINT_TYPE_PRESOLVER_SOLUTION) type = IterationType::kPresolveTermination;
  if (callback_) callback_(IterationCallbackInfo{type, params.termination_criteria.get(), log.solution_stats.get(), original_bound_norms_});
  if (Logs(params, 1)) {  // :1617-1641
    LogInfoWithoutPrefix(std::string("Termination reason: ") + TerminationReasonName(log.termination_reason));
    LogInfoWithoutPrefix(std::string("Solution point type: ") + PointTypeName(log.solution_type));
    LogInfoWithoutPrefix("Final solution stats:");
    LogInfoWithoutPrefix(IterationStatsHeader(params.verbosity_level, params.use_feasibility_polishing));
    LogInfoWithoutPrefix(IterationStatsLine(params.verbosity_level, params.use_feasibility_polishing, type, log.solution_stats.get(), params.termination_criteria.get(), original_bound_norms_, log.solution_type));
    const auto info = GetConvergenceInformationFor(log.solution_stats.get(), log.solution_type);
    if (info.has_value() && std::isfinite(info->corrected_dual_objective.get())) {
      LogInfoWithoutPrefix(Format("Dual objective after infeasibility correction: %g", info->corrected_dual_objective.get()));
    }
  }
  result.solve_log = std::move(log);
  return result;
}

}  // namespace pdlp_b200
