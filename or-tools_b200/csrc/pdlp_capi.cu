// The C++ entry points (primal_dual_hybrid_gradient.h) and the C ABI (include/pdlp_b200.h).
//   follows ortools/pdlp/primal_dual_hybrid_gradient.cc:2778-2831 (entry checks),
//           ortools/pdlp/quadratic_program.cc:37-90 (dimension / bound validation).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "pdlp_b200.h"
#include "pdlp_driver.h"
#include "pdlp_logging.h"

namespace pdlp_b200 {

std::string ValidateQuadraticProgramDimensions(const QuadraticProgram& qp) {  // quadratic_program.cc:37-80
  const int64_t var_lb_size = qp.variable_lower_bounds.size();
  const int64_t con_lb_size = qp.constraint_lower_bounds.size();
  auto S = [](int64_t v) { return std::to_string(v); };
  if (var_lb_size != static_cast<int64_t>(qp.variable_upper_bounds.size()))
    return "Inconsistent dimensions: variable lower bound vector has size " + S(var_lb_size) + " while variable upper bound vector has size " +
           S(qp.variable_upper_bounds.size());
  if (var_lb_size != static_cast<int64_t>(qp.objective_vector.size()))
    return "Inconsistent dimensions: variable lower bound vector has size " + S(var_lb_size) + " while objective vector has size " +
           S(qp.objective_vector.size());
  if (var_lb_size != qp.constraint_matrix.cols)
    return "Inconsistent dimensions: variable lower bound vector has size " + S(var_lb_size) + " while constraint matrix has " +
           S(qp.constraint_matrix.cols) + " columns";
  if (qp.objective_matrix.has_value() && var_lb_size != static_cast<int64_t>(qp.objective_matrix->size()))
    return "Inconsistent dimensions: variable lower bound vector has size " + S(var_lb_size) + " while objective matrix has " +
           S(qp.objective_matrix->size()) + " rows";
  if (con_lb_size != static_cast<int64_t>(qp.constraint_upper_bounds.size()))
    return "Inconsistent dimensions: constraint lower bound vector has size " + S(con_lb_size) + " while constraint upper bound vector has size " +
           S(qp.constraint_upper_bounds.size());
  if (con_lb_size != qp.constraint_matrix.rows)
    return "Inconsistent dimensions: constraint lower bound vector has size " + S(con_lb_size) + " while constraint matrix has " +
           S(qp.constraint_matrix.rows) + " rows ";
  return "";
}

namespace {

// Runs pred over [0, count) in slices on several host threads (the data checks scan every bound
// and diagonal entry once: 1 GB of reads on the 1e8-nonzero LP).  True iff every slice is true.
template <class SlicePred>
bool AllSlices(int64_t count, SlicePred pred) {
  constexpr int64_t kSerialBelow = int64_t{1} << 20;
  const unsigned hw = std::thread::hardware_concurrency();
  const int threads = count < kSerialBelow ? 1 : static_cast<int>(std::max(1u, std::min(8u, hw / 2)));
  if (threads == 1) return pred(int64_t{0}, count);
  std::vector<char> ok(threads, 1);
  std::vector<std::thread> pool;
  const int64_t per = (count + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    pool.emplace_back([&, t]() {
      const int64_t lo = std::min<int64_t>(count, t * per), hi = std::min<int64_t>(count, lo + per);
      ok[t] = pred(lo, hi) ? 1 : 0;
    });
  }
  for (std::thread& th : pool) th.join();
  for (char c : ok) if (!c) return false;
  return true;
}

bool BoundsAreValid(const double* lb, const double* ub, int64_t count) {
  return AllSlices(count, [&](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) if (!(lb[i] <= ub[i])) return false;
    return true;
  });
}

// The checks of pdhg.cc:2787-2826 that need the data (not the dimensions), then the solve.
SolverResult SolveView(const QpView& qp, const PrimalDualHybridGradientParams& params, const double* x0, const double* y0,
                       const InterruptFlag* interrupt, IterationCallback callback, SolutionSink* sink = nullptr) {
  // Data checks in the reference's order; the first that fails wins (smallest code).
  auto data_error = [&]() -> int {
    if (!BoundsAreValid(qp.clb, qp.cub, qp.m) || !BoundsAreValid(qp.vlb, qp.vub, qp.n)) return 3;
    if (qp.objective_scaling_factor == 0) return 2;
    if (qp.q != nullptr) {
      const bool convex = AllSlices(qp.n, [&](int64_t lo, int64_t hi) {
        for (int64_t j = lo; j < hi; ++j) if (qp.q[j] < 0.0) return false;
        return true;
      });
      if (!convex) return 1;
    }
    return 0;
  };
  auto describe = [](int code) {
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM,
                             code == 3   ? "The input problem has inconsistent bounds."
                             : code == 2 ? "The objective scaling factor cannot be zero."
                                         : "The objective is not convex (i.e., the objective matrix contains negative entries).");
  };
  int code = data_error();
  if (qp.world <= 1 && code != 0) return describe(code);
  if (params.use_feasibility_polishing && qp.q != nullptr)
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, "use_feasibility_polishing is only implemented for linear programs.");
  if (params.presolve_options->use_glop)
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER,
                             "presolve_options.use_glop is not supported by the device path: Glop presolve is host-only. Presolve on the host and call pdlp_b200_solve_presolved with the presolved problem and a postsolve callback.");
  WallTimer upload;
  upload.Start();
  PreprocessSolver solver(qp, params);
  solver.set_solution_sink(sink);
  // Sharded: a rank whose block is fine must fail with the rank whose block is not.
  if (qp.world > 1 && (code = solver.AgreeOnCode(code)) != 0) return describe(code);
  const double upload_seconds = upload.Get();
  SolverResult result = solver.PreprocessAndSolve(params, x0, y0, interrupt, std::move(callback));
  g_last_counters.upload_seconds = upload_seconds;
  return result;
}

// The presolve hand-off (pdhg.cc:1145-1202, :986-1034, :1288-1341 with the caller's presolver in Glop's place):
// PDHG runs on `presolved` (null: the presolver solved the problem, `presolve_reason` says how), every
// statistic, termination test and the result are those of `original` at the point `postsolve` recovers.
SolverResult SolveViewPresolved(const QpView& original, const QpView* presolved, int presolve_reason,
                                std::function<void(const double*, const double*, double*, double*)> postsolve, const PrimalDualHybridGradientParams& params,
                                const double* x0, const double* y0, const InterruptFlag* interrupt, IterationCallback callback, SolutionSink* sink) {
  auto invalid = [&](const QpView& qp) -> const char* {
    if (!BoundsAreValid(qp.clb, qp.cub, qp.m) || !BoundsAreValid(qp.vlb, qp.vub, qp.n)) return "The input problem has inconsistent bounds.";
    if (qp.objective_scaling_factor == 0) return "The objective scaling factor cannot be zero.";
    if (qp.q != nullptr) {
      for (int64_t j = 0; j < qp.n; ++j) if (qp.q[j] < 0.0) return "The objective is not convex (i.e., the objective matrix contains negative entries).";
    }
    return nullptr;
  };
  if (const char* e = invalid(original)) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, e);
  if (presolved != nullptr) {
    if (const char* e = invalid(*presolved)) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, std::string("presolved problem: ") + e);
  }
  if (params.use_feasibility_polishing && (original.q != nullptr || (presolved != nullptr && presolved->q != nullptr)))
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, "use_feasibility_polishing is only implemented for linear programs.");
  WallTimer upload;
  upload.Start();
  PreprocessSolver original_solver(original, params);
  PresolveHook hook;
  hook.original = &original_solver;
  hook.postsolve = std::move(postsolve);
  static const int64_t kNoColumns[1] = {0};
  QpView empty;  // zero variables, zero constraints: what PDHG "iterates on" when the presolver solved everything
  empty.col_ptr = kNoColumns;
  PreprocessSolver working(presolved != nullptr ? *presolved : empty, params);
  working.set_presolve(&hook);
  working.set_solution_sink(sink);
  const double upload_seconds = upload.Get();
  SolverResult result = presolved != nullptr ? working.PreprocessAndSolve(params, x0, y0, interrupt, std::move(callback))
                                             : working.FinishPresolvedOnly(params, presolve_reason, interrupt, std::move(callback));
  g_last_counters.upload_seconds = upload_seconds;
  return result;
}

QpView ViewOf(const QuadraticProgram& qp) {
  QpView v;
  v.n = qp.variable_lower_bounds.size();
  v.m = qp.constraint_lower_bounds.size();
  v.nnz = qp.constraint_matrix.nonZeros();
  v.col_ptr = qp.constraint_matrix.outer_index.data();
  v.row_idx = qp.constraint_matrix.inner_index.data();
  v.val = qp.constraint_matrix.values.data();
  v.c = qp.objective_vector.data();
  v.q = qp.objective_matrix.has_value() ? qp.objective_matrix->data() : nullptr;
  v.vlb = qp.variable_lower_bounds.data(); v.vub = qp.variable_upper_bounds.data();
  v.clb = qp.constraint_lower_bounds.data(); v.cub = qp.constraint_upper_bounds.data();
  v.objective_offset = qp.objective_offset;
  v.objective_scaling_factor = qp.objective_scaling_factor;
  v.name = qp.problem_name.has_value() ? qp.problem_name->c_str() : nullptr;
  return v;
}

}  // namespace

bool HasValidBounds(const QuadraticProgram& qp) {  // quadratic_program.cc:82-90
  return BoundsAreValid(qp.constraint_lower_bounds.data(), qp.constraint_upper_bounds.data(), qp.constraint_lower_bounds.size()) &&
         BoundsAreValid(qp.variable_lower_bounds.data(), qp.variable_upper_bounds.data(), qp.variable_lower_bounds.size());
}

SolverResult PrimalDualHybridGradient(QuadraticProgram qp, const PrimalDualHybridGradientParams& params, const std::atomic<bool>* interrupt_solve,
                                      std::function<void(const IterationCallbackInfo&)> iteration_stats_callback) {
  return PrimalDualHybridGradient(std::move(qp), params, std::nullopt, interrupt_solve, std::move(iteration_stats_callback));
}

SolverResult PrimalDualHybridGradient(QuadraticProgram qp, const PrimalDualHybridGradientParams& params,
                                      std::optional<PrimalAndDualSolution> initial_solution, const std::atomic<bool>* interrupt_solve,
                                      std::function<void(const IterationCallbackInfo&)> iteration_stats_callback) {
  const std::string perr = ValidatePrimalDualHybridGradientParams(params);
  if (!perr.empty()) return ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, perr);
  const int64_t n = qp.variable_lower_bounds.size();
  if (static_cast<int64_t>(qp.constraint_matrix.outer_index.size()) != qp.constraint_matrix.cols + 1 ||
      qp.constraint_matrix.inner_index.size() != qp.constraint_matrix.values.size()) {
    return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "constraint_matrix must be in compressed format. Call constraint_matrix.makeCompressed()");
  }
  const std::string derr = ValidateQuadraticProgramDimensions(qp);
  if (!derr.empty()) return ErrorSolverResult(TERMINATION_REASON_INVALID_PROBLEM, "INVALID_ARGUMENT: " + derr);
  const int64_t m = qp.constraint_lower_bounds.size();
  const double* x0 = nullptr;
  const double* y0 = nullptr;
  if (initial_solution.has_value()) {  // size checks of pdhg.cc:899-947
    if (static_cast<int64_t>(initial_solution->primal_solution.size()) != n)
      return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION, "Initial primal solution has size " + std::to_string(initial_solution->primal_solution.size()) +
                                                                                " which differs from problem primal size " + std::to_string(n));
    if (static_cast<int64_t>(initial_solution->dual_solution.size()) != m)
      return ErrorSolverResult(TERMINATION_REASON_INVALID_INITIAL_SOLUTION, "Initial dual solution has size " + std::to_string(initial_solution->dual_solution.size()) +
                                                                                " which differs from problem dual size " + std::to_string(m));
    x0 = initial_solution->primal_solution.data();
    y0 = initial_solution->dual_solution.data();
    // A zero-sized vector's data() may be null: keep "given" distinguishable from "absent".
    static const double kEmpty = 0.0;
    if (x0 == nullptr) x0 = &kEmpty;
    if (y0 == nullptr) y0 = &kEmpty;
  }
  InterruptFlag flag;
  flag.atomic_flag = interrupt_solve;
  return SolveView(ViewOf(qp), params, x0, y0, interrupt_solve != nullptr ? &flag : nullptr, std::move(iteration_stats_callback));
}

}  // namespace pdlp_b200

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
using namespace pdlp_b200;

namespace {

thread_local std::string g_last_error;

pdlp_b200_status Fail(pdlp_b200_status s, const std::string& msg) {
  g_last_error = msg;
  return s;
}

template <class F>
pdlp_b200_status Guard(F&& f) {
  try {
    return f();
  } catch (const std::invalid_argument& e) {
    return Fail(PDLP_B200_BAD_ARGUMENT, e.what());
  } catch (const std::runtime_error& e) {
    const std::string what = e.what();
    return Fail(what.rfind("CUDA error", 0) == 0 ? PDLP_B200_CUDA_ERROR : PDLP_B200_INTERNAL_ERROR, what);
  } catch (const std::exception& e) {
    return Fail(PDLP_B200_INTERNAL_ERROR, e.what());
  }
}

void SelectDevice(int device) {
  int count = 0;
  PDLP_CUDA_CHECK(cudaGetDeviceCount(&count));
  if (count == 0) throw std::runtime_error("CUDA error: no CUDA device (pdlp_b200 has no CPU fallback)");
  if (device >= 0) PDLP_CUDA_CHECK(cudaSetDevice(device));
}

uint8_t* DupBytes(const std::string& s) {
  uint8_t* p = static_cast<uint8_t*>(std::malloc(s.size() > 0 ? s.size() : 1));
  if (!s.empty()) std::memcpy(p, s.data(), s.size());
  return p;
}
char* DupString(const std::string& s) {
  char* p = static_cast<char*>(std::malloc(s.size() + 1));
  std::memcpy(p, s.c_str(), s.size() + 1);
  return p;
}

bool ParseParams(const uint8_t* bytes, size_t len, PrimalDualHybridGradientParams* out) {
  if (bytes == nullptr || len == 0) return true;
  return Parse(bytes, len, out);
}

bool ProblemPointersOk(const pdlp_b200_problem* p) {
  if (p == nullptr || p->num_variables < 0 || p->num_constraints < 0 || p->num_nonzeros < 0) return false;
  if (p->col_ptr == nullptr && p->col_ptr32 == nullptr) return false;
  if (p->num_nonzeros > 0 && p->values == nullptr) return false;
  if (p->num_nonzeros > 0 && (p->col_ptr32 != nullptr ? p->row_idx32 == nullptr : p->row_idx == nullptr)) return false;
  if (p->num_variables > 0 && (p->objective_vector == nullptr || p->variable_lower_bounds == nullptr || p->variable_upper_bounds == nullptr)) return false;
  if (p->num_constraints > 0 && (p->constraint_lower_bounds == nullptr || p->constraint_upper_bounds == nullptr)) return false;
  return true;
}

QpView ViewOfC(const pdlp_b200_problem* p) {
  QpView v;
  v.n = p->num_variables; v.m = p->num_constraints; v.nnz = p->num_nonzeros;
  v.col_ptr = p->col_ptr; v.row_idx = p->row_idx; v.val = p->values;
  v.col_ptr32 = p->col_ptr32; v.row_idx32 = p->row_idx32;
  v.c = p->objective_vector; v.q = p->objective_matrix_diagonal;
  v.vlb = p->variable_lower_bounds; v.vub = p->variable_upper_bounds;
  v.clb = p->constraint_lower_bounds; v.cub = p->constraint_upper_bounds;
  v.objective_offset = p->objective_offset; v.objective_scaling_factor = p->objective_scaling_factor;
  v.name = p->problem_name;
  return v;
}

// Adds the shard description; returns an error text when it is unusable.
std::string ApplyShard(const pdlp_b200_shard* s, QpView* v) {
  if (s == nullptr) return "";
  if (s->world < 1 || s->rank < 0 || s->rank >= s->world) return "shard: rank / world out of range";
  if (s->world == 1) return "";
  if (s->nccl_unique_id == nullptr) return "shard: nccl_unique_id is null";
  if (s->column_offset < 0 || s->global_num_variables < s->column_offset + v->n) return "shard: column block lies outside the global problem";
  if (v->n < 1 || v->m < 1) return "shard: every rank needs at least one column, and the problem one constraint";
  v->global_n = s->global_num_variables;
  v->col_offset = s->column_offset;
  v->rank = s->rank;
  v->world = s->world;
  v->nccl_unique_id = s->nccl_unique_id;
  if (s->num_shared_constraints > v->m) return "shard: num_shared_constraints exceeds the block's constraint count";
  v->num_shared_rows = s->num_shared_constraints < 0 ? v->m : s->num_shared_constraints;
  return "";
}

}  // namespace

struct pdlp_b200_device_problem {
  std::unique_ptr<PreprocessSolver> pre;
  std::unique_ptr<Iterate> iterate;
  int device = 0;
};

extern "C" {

const char* pdlp_b200_last_error(void) { return g_last_error.c_str(); }
void pdlp_b200_free(void* p) { std::free(p); }

pdlp_b200_status pdlp_b200_solve(const pdlp_b200_problem* problem, const uint8_t* params_bytes, size_t params_len, const double* x0, const double* y0,
                                 const volatile int* interrupt, pdlp_b200_callback callback, void* callback_user, int device, double* primal,
                                 double* dual, double* reduced, int* has_solution, uint8_t** solve_log, size_t* solve_log_len,
                                 pdlp_b200_counters* counters) {
  return pdlp_b200_solve_sharded(problem, nullptr, params_bytes, params_len, x0, y0, interrupt, callback, callback_user, device, primal, dual, reduced,
                                 has_solution, solve_log, solve_log_len, counters);
}

pdlp_b200_status pdlp_b200_unique_id(uint8_t* out) {
  return Guard([&]() -> pdlp_b200_status {
    if (out == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "out is null");
    NcclApi& api = NcclApi::Get();
    if (!api.ok()) return Fail(PDLP_B200_INTERNAL_ERROR, "NCCL could not be loaded: " + api.error());
    ncclUniqueId id;
    const ncclResult_t r = api.GetUniqueId(&id);
    if (r != ncclSuccess) return Fail(PDLP_B200_INTERNAL_ERROR, std::string("ncclGetUniqueId failed: ") + api.GetErrorString(r));
    static_assert(NCCL_UNIQUE_ID_BYTES == PDLP_B200_UNIQUE_ID_BYTES, "unique id size");
    std::memcpy(out, id.internal, NCCL_UNIQUE_ID_BYTES);
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_release_communicators(int* released) {
  return Guard([&]() -> pdlp_b200_status {
    const int count = CommCache::Get().Release();
    if (released != nullptr) *released = count;
    return PDLP_B200_OK;
  });
}

namespace {
struct PresolveArgs {
  const pdlp_b200_problem* presolved;  // null: the presolver solved the problem
  int presolve_reason;
  pdlp_b200_postsolve postsolve;
  void* user;
};
}  // namespace

static pdlp_b200_status SolveCommon(const pdlp_b200_problem* problem, const pdlp_b200_shard* shard, const PresolveArgs* pre, const uint8_t* params_bytes,
                                    size_t params_len, const double* x0, const double* y0, const volatile int* interrupt, pdlp_b200_callback callback,
                                    void* callback_user, int device, double* primal, double* dual, double* reduced, int* has_solution, uint8_t** solve_log,
                                    size_t* solve_log_len, pdlp_b200_counters* counters) {
  return Guard([&]() -> pdlp_b200_status {
    if (!ProblemPointersOk(problem) || solve_log == nullptr || solve_log_len == nullptr || has_solution == nullptr)
      return Fail(PDLP_B200_BAD_ARGUMENT, "null or negative-sized argument");
    if (pre != nullptr) {
      if (pre->postsolve == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "postsolve callback is null");
      if (pre->presolved != nullptr && !ProblemPointersOk(pre->presolved)) return Fail(PDLP_B200_BAD_ARGUMENT, "presolved problem: null or negative-sized argument");
      if (pre->presolved == nullptr && (pre->presolve_reason < TERMINATION_REASON_OPTIMAL || pre->presolve_reason > TERMINATION_REASON_INVALID_INITIAL_SOLUTION))
        return Fail(PDLP_B200_BAD_ARGUMENT, "presolve_status must be a TerminationReason when there is no presolved problem");
    }
    QpView view = ViewOfC(problem);
    if (const std::string e = ApplyShard(shard, &view); !e.empty()) return Fail(PDLP_B200_BAD_ARGUMENT, e);
    if ((x0 == nullptr) != (y0 == nullptr)) {
      // A zero-length side may legitimately be null when the other is given.
      if (!((x0 == nullptr && problem->num_variables == 0) || (y0 == nullptr && problem->num_constraints == 0)))
        return Fail(PDLP_B200_BAD_ARGUMENT, "initial primal and dual solutions must be given together");
    }
    PrimalDualHybridGradientParams params;
    if (!ParseParams(params_bytes, params_len, &params)) return Fail(PDLP_B200_BAD_ARGUMENT, "params is not a serialized PrimalDualHybridGradientParams");
    *has_solution = 0;
    SolverResult result;
    SolutionSink sink;
    sink.primal = primal; sink.dual = dual; sink.reduced_costs = reduced;
    const std::string perr = ValidatePrimalDualHybridGradientParams(params);
    if (!perr.empty()) {
      result = ErrorSolverResult(TERMINATION_REASON_INVALID_PARAMETER, perr);
    } else {
      SelectDevice(device);
      InterruptFlag flag;
      flag.int_flag = interrupt;
      IterationCallback cb;
      if (callback != nullptr) {
        cb = [&](const IterationCallbackInfo& info) {
          const std::string bytes = Serialize(info.iteration_stats);
          const double norms[4] = {info.bound_norms.l2_norm_primal_linear_objective, info.bound_norms.l2_norm_constraint_bounds,
                                   info.bound_norms.l_inf_norm_primal_linear_objective, info.bound_norms.l_inf_norm_constraint_bounds};
          callback(static_cast<int>(info.iteration_type), reinterpret_cast<const uint8_t*>(bytes.data()), bytes.size(), norms, callback_user);
        };
      }
      static const double kEmpty = 0.0;
      const bool warm = x0 != nullptr || y0 != nullptr;
      if (pre != nullptr) {
        QpView presolved_view;
        if (pre->presolved != nullptr) presolved_view = ViewOfC(pre->presolved);
        pdlp_b200_postsolve fn = pre->postsolve;
        void* user = pre->user;
        result = SolveViewPresolved(view, pre->presolved != nullptr ? &presolved_view : nullptr, pre->presolve_reason,
                                    [fn, user](const double* px, const double* py, double* ox, double* oy) { fn(px, py, ox, oy, user); }, params,
                                    warm ? (x0 ? x0 : &kEmpty) : nullptr, warm ? (y0 ? y0 : &kEmpty) : nullptr, interrupt != nullptr ? &flag : nullptr, std::move(cb),
                                    &sink);
      } else {
        result = SolveView(view, params, warm ? (x0 ? x0 : &kEmpty) : nullptr, warm ? (y0 ? y0 : &kEmpty) : nullptr,
                           interrupt != nullptr ? &flag : nullptr, std::move(cb), &sink);
      }
    }
    // The solver wrote the solution straight into primal / dual / reduced (SolutionSink).
    if (sink.filled && result.solve_log.termination_reason != TERMINATION_REASON_INVALID_PROBLEM &&
        result.solve_log.termination_reason != TERMINATION_REASON_INVALID_PARAMETER &&
        result.solve_log.termination_reason != TERMINATION_REASON_INVALID_INITIAL_SOLUTION) {
      *has_solution = 1;
    }
    const std::string bytes = Serialize(result.solve_log);
    *solve_log = DupBytes(bytes);
    *solve_log_len = bytes.size();
    if (counters != nullptr) {
      const SolveCounters& c = LastSolveCounters();
      counters->kernel_launches = c.kernel_launches;
      counters->trial_steps = c.trial_steps;
      counters->setup_seconds = c.setup_seconds;
      counters->loop_seconds = c.loop_seconds;
      counters->device_loop_seconds = c.device_loop_seconds;
      counters->collectives = c.collectives;
      counters->upload_seconds = c.upload_seconds;
      counters->symmetric_buffers = c.symmetric_buffers;
      counters->peer_exchanges = c.peer_exchanges;
      counters->interrupts_seen_on_device = c.interrupts_seen_on_device;
    }
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_solve_sharded(const pdlp_b200_problem* problem, const pdlp_b200_shard* shard, const uint8_t* params_bytes, size_t params_len,
                                         const double* x0, const double* y0, const volatile int* interrupt, pdlp_b200_callback callback,
                                         void* callback_user, int device, double* primal, double* dual, double* reduced, int* has_solution,
                                         uint8_t** solve_log, size_t* solve_log_len, pdlp_b200_counters* counters) {
  return SolveCommon(problem, shard, nullptr, params_bytes, params_len, x0, y0, interrupt, callback, callback_user, device, primal, dual, reduced, has_solution, solve_log,
                     solve_log_len, counters);
}

pdlp_b200_status pdlp_b200_solve_presolved(const pdlp_b200_problem* original, const pdlp_b200_problem* presolved, int presolve_status, pdlp_b200_postsolve postsolve,
                                           void* postsolve_user, const uint8_t* params_bytes, size_t params_len, const double* x0, const double* y0,
                                           const volatile int* interrupt, pdlp_b200_callback callback, void* callback_user, int device, double* primal, double* dual,
                                           double* reduced, int* has_solution, uint8_t** solve_log, size_t* solve_log_len, pdlp_b200_counters* counters) {
  const PresolveArgs pre{presolved, presolve_status, postsolve, postsolve_user};
  return SolveCommon(original, nullptr, &pre, params_bytes, params_len, x0, y0, interrupt, callback, callback_user, device, primal, dual, reduced, has_solution,
                     solve_log, solve_log_len, counters);
}

pdlp_b200_status pdlp_b200_validate_params(const uint8_t* params_bytes, size_t params_len, char** message) {
  return Guard([&]() -> pdlp_b200_status {
    if (message == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "message is null");
    PrimalDualHybridGradientParams params;
    if (!ParseParams(params_bytes, params_len, &params)) return Fail(PDLP_B200_BAD_ARGUMENT, "params is not a serialized PrimalDualHybridGradientParams");
    const std::string e = ValidatePrimalDualHybridGradientParams(params);
    *message = e.empty() ? nullptr : DupString(e);
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_check_termination(const uint8_t* params_bytes, size_t params_len, const uint8_t* stats_bytes, size_t stats_len,
                                             const double* bound_norms, int which, int force_numerical_termination, const volatile int* interrupt,
                                             int* reason, int* point_type, double* relative_residuals) {
  return Guard([&]() -> pdlp_b200_status {
    if (reason == nullptr || point_type == nullptr || bound_norms == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "null argument");
    PrimalDualHybridGradientParams params;
    if (!ParseParams(params_bytes, params_len, &params)) return Fail(PDLP_B200_BAD_ARGUMENT, "params is not a serialized PrimalDualHybridGradientParams");
    IterationStats stats;
    if (stats_len > 0 && !Parse(stats_bytes, stats_len, &stats)) return Fail(PDLP_B200_BAD_ARGUMENT, "stats is not a serialized IterationStats");
    const TerminationCriteria& criteria = params.termination_criteria.get();
    const QuadraticProgramBoundNorms norms{bound_norms[0], bound_norms[1], bound_norms[2], bound_norms[3]};
    std::optional<TerminationReasonAndPointType> r;
    if (which == 0) {
      r = CheckIterateTerminationCriteria(criteria, stats, norms, force_numerical_termination != 0);
    } else {
      InterruptFlag flag;
      flag.int_flag = interrupt;
      r = CheckSimpleTerminationCriteria(criteria, stats, &flag);
    }
    *reason = r.has_value() ? r->reason : TERMINATION_REASON_UNSPECIFIED;
    *point_type = r.has_value() ? r->type : POINT_TYPE_UNSPECIFIED;
    if (relative_residuals != nullptr) {
      for (int k = 0; k < 5; ++k) relative_residuals[k] = 0.0;
      if (!stats.convergence_information.empty()) {
        const RelativeConvergenceInformation rel = ComputeRelativeResiduals(EffectiveOptimalityCriteria(criteria), stats.convergence_information[0], norms);
        relative_residuals[0] = rel.relative_l_inf_primal_residual; relative_residuals[1] = rel.relative_l2_primal_residual;
        relative_residuals[2] = rel.relative_l_inf_dual_residual; relative_residuals[3] = rel.relative_l2_dual_residual;
        relative_residuals[4] = rel.relative_optimality_gap;
      }
    }
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_format_iteration_stats(const uint8_t* params, size_t params_len, const uint8_t* stats, size_t stats_len, const double* bound_norms,
                                                  int iteration_type, int preferred_candidate, char** header, char** line) {
  return Guard([&]() -> pdlp_b200_status {
    if (stats == nullptr || bound_norms == nullptr || header == nullptr || line == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PrimalDualHybridGradientParams p;
    if (!ParseParams(params, params_len, &p)) return Fail(PDLP_B200_BAD_ARGUMENT, "params bytes do not parse as PrimalDualHybridGradientParams");
    IterationStats s;
    if (!Parse(stats, stats_len, &s)) return Fail(PDLP_B200_BAD_ARGUMENT, "stats bytes do not parse as IterationStats");
    QuadraticProgramBoundNorms bn;
    bn.l2_norm_primal_linear_objective = bound_norms[0]; bn.l2_norm_constraint_bounds = bound_norms[1];
    bn.l_inf_norm_primal_linear_objective = bound_norms[2]; bn.l_inf_norm_constraint_bounds = bound_norms[3];
    *header = DupString(IterationStatsHeader(p.verbosity_level, p.use_feasibility_polishing));
    *line = DupString(IterationStatsLine(p.verbosity_level, p.use_feasibility_polishing, static_cast<IterationType>(iteration_type), s, p.termination_criteria.get(), bn,
                                         preferred_candidate));
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_default_params(uint8_t** params, size_t* params_len) {
  return Guard([&]() -> pdlp_b200_status {
    if (params == nullptr || params_len == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "null output");
    const std::string bytes = Serialize(PrimalDualHybridGradientParams());
    *params = DupBytes(bytes);
    *params_len = bytes.size();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_describe_schema(char** text) {
  return Guard([&]() -> pdlp_b200_status {
    if (text == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "null output");
    *text = DupString(DescribeSchema());
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_problem_create(const pdlp_b200_problem* problem, int device, pdlp_b200_device_problem** out) {
  return pdlp_b200_problem_create_sharded(problem, nullptr, device, out);
}

pdlp_b200_status pdlp_b200_problem_create_sharded(const pdlp_b200_problem* problem, const pdlp_b200_shard* shard, int device,
                                                  pdlp_b200_device_problem** out) {
  return Guard([&]() -> pdlp_b200_status {
    if (!ProblemPointersOk(problem) || out == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "null or negative-sized argument");
    QpView v = ViewOfC(problem);
    if (const std::string e = ApplyShard(shard, &v); !e.empty()) return Fail(PDLP_B200_BAD_ARGUMENT, e);
    SelectDevice(device);
    auto h = std::make_unique<pdlp_b200_device_problem>();
    PDLP_CUDA_CHECK(cudaGetDevice(&h->device));
    v.name = nullptr;
    h->pre = std::make_unique<PreprocessSolver>(v, PrimalDualHybridGradientParams());
    *out = h.release();
    return PDLP_B200_OK;
  });
}

void pdlp_b200_problem_destroy(pdlp_b200_device_problem* h) {
  if (h == nullptr) return;
  cudaSetDevice(h->device);
  h->iterate.reset();
  h->pre.reset();
  delete h;
}

pdlp_b200_status pdlp_b200_compute_statuses(const pdlp_b200_problem* problem, const double* primal_solution, const double* dual_solution,
                                            int8_t* variable_statuses, int8_t* constraint_statuses) {
  return Guard([&]() -> pdlp_b200_status {
    if (problem == nullptr || problem->num_variables < 0 || problem->num_constraints < 0) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    const int64_t n = problem->num_variables, m = problem->num_constraints;
    if ((n > 0 && (primal_solution == nullptr || variable_statuses == nullptr || problem->variable_lower_bounds == nullptr || problem->variable_upper_bounds == nullptr)) ||
        (m > 0 && (dual_solution == nullptr || constraint_statuses == nullptr || problem->constraint_lower_bounds == nullptr || problem->constraint_upper_bounds == nullptr))) {
      return Fail(PDLP_B200_BAD_ARGUMENT, "null vector");
    }
    // The same rules as internal::ComputeStatuses of the C++ mirror (include/pdlp_b200/primal_dual_hybrid_gradient.h), on raw arrays.
    for (int64_t i = 0; i < m; ++i) {
      ConstraintStatus s = ConstraintStatus::BASIC;
      if (problem->constraint_lower_bounds[i] == problem->constraint_upper_bounds[i]) s = ConstraintStatus::FIXED_VALUE;
      else if (dual_solution[i] > 0) s = ConstraintStatus::AT_LOWER_BOUND;
      else if (dual_solution[i] < 0) s = ConstraintStatus::AT_UPPER_BOUND;
      constraint_statuses[i] = static_cast<int8_t>(s);
    }
    for (int64_t j = 0; j < n; ++j) {
      const bool at_lb = primal_solution[j] <= problem->variable_lower_bounds[j], at_ub = primal_solution[j] >= problem->variable_upper_bounds[j];
      const VariableStatus s = at_lb ? (at_ub ? VariableStatus::FIXED_VALUE : VariableStatus::AT_LOWER_BOUND) : (at_ub ? VariableStatus::AT_UPPER_BOUND : VariableStatus::BASIC);
      variable_statuses[j] = static_cast<int8_t>(s);
    }
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_work_blocks(const int64_t* offsets, int64_t rows, int64_t cut, int32_t* block_start, int64_t* num_blocks,
                                       int32_t* cut_block, int32_t* max_row_nnz) {
  return Guard([&]() -> pdlp_b200_status {
    if (offsets == nullptr || block_start == nullptr || num_blocks == nullptr || rows < 0 || rows >= (int64_t{1} << 31) - 1 || cut >= rows || cut == 0) {
      return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    }
    for (int64_t r = 0; r < rows; ++r) {
      if (offsets[r + 1] < offsets[r]) return Fail(PDLP_B200_BAD_ARGUMENT, "offsets must not decrease");
    }
    if (offsets[rows] - offsets[0] >= (int64_t{1} << 31) - 1) return Fail(PDLP_B200_BAD_ARGUMENT, "too many nonzeros for one device block");
    int mx = 0, cb = -1;
    const std::vector<int> starts = PartitionRowsWithCut(offsets, static_cast<int>(rows), cut < 0 ? -1 : static_cast<int>(cut), &mx, &cb);
    std::copy(starts.begin(), starts.end(), block_start);
    *num_blocks = static_cast<int64_t>(starts.size()) - 1;
    if (cut_block != nullptr) *cut_block = cb;
    if (max_row_nnz != nullptr) *max_row_nnz = mx;
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_matvec(pdlp_b200_device_problem* h, int which, const double* v, double* out, int repeat, double* seconds_per_launch) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || v == nullptr || out == nullptr || (which != 0 && which != 1)) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    const int64_t in_size = which == 0 ? p.m : p.n, out_size = which == 0 ? p.n : p.m;
    DVec& din = which == 0 ? ws.scratch_m[0] : ws.scratch_n[0];
    DVec& dout = which == 0 ? ws.scratch_n[1] : ws.scratch_m[1];
    din.Upload(v, in_size, ws.stream);
    const SparseRows& s = which == 0 ? p.K.by_col : p.K.by_row;
    if (repeat < 1) repeat = 1;
    cudaEvent_t e0, e1;
    PDLP_CUDA_CHECK(cudaEventCreate(&e0));
    PDLP_CUDA_CHECK(cudaEventCreate(&e1));
    RowDot(ws, s, din.get(), dout.get());  // warm-up (also the result)
    PDLP_CUDA_CHECK(cudaEventRecord(e0, ws.stream));
    for (int r = 0; r < repeat; ++r) RowDot(ws, s, din.get(), dout.get());
    PDLP_CUDA_CHECK(cudaEventRecord(e1, ws.stream));
    dout.Download(out, out_size, ws.stream);
    PDLP_CUDA_CHECK(cudaGetLastError());
    ws.Sync();
    float ms = 0;
    PDLP_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (seconds_per_launch != nullptr) *seconds_per_launch = ms * 1e-3 / repeat;
    return PDLP_B200_OK;
  });
}

// No counterpart in the reference: K v from caller-owned device arrays with the
// work-block table in a static device array; kernels and scratch in pdlp_noalloc.cu.
pdlp_b200_status pdlp_b200_matvec_no_alloc(int64_t rows, int64_t cols, const int32_t* row_ptr, const int32_t* col_idx, const double* values, const double* v,
                                           double* out, int device, void* stream) {
  return Guard([&]() -> pdlp_b200_status {
    if (rows < 0 || cols < 0 || row_ptr == nullptr || (rows > 0 && out == nullptr)) return Fail(PDLP_B200_BAD_ARGUMENT, "pdlp_b200_matvec_no_alloc: bad argument");
    if (rows > PDLP_B200_NO_ALLOC_MAX_ROWS) {
      return Fail(PDLP_B200_BAD_ARGUMENT, "pdlp_b200_matvec_no_alloc: " + std::to_string(rows) + " rows exceed the static scratch limit of " +
                                              std::to_string(PDLP_B200_NO_ALLOC_MAX_ROWS) + " rows (use pdlp_b200_problem_create + pdlp_b200_matvec)");
    }
    SelectDevice(device);
    MatVecNoAlloc(rows, cols, row_ptr, col_idx, values, v, out, static_cast<cudaStream_t>(stream));
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_scaled_norms(pdlp_b200_device_problem* h, int which, int l2, const double* row_scale, const double* col_scale, double* out) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || row_scale == nullptr || col_scale == nullptr || out == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    ws.scratch_m[0].Upload(row_scale, p.m, ws.stream);
    ws.scratch_n[0].Upload(col_scale, p.n, ws.stream);
    if (which == 0) {
      ScaledRowNorms(ws, p.K.by_col, ws.scratch_m[0].get(), ws.scratch_n[0].get(), l2 != 0, ws.scratch_n[1].get());
      ws.scratch_n[1].Download(out, p.n, ws.stream);
    } else {
      ScaledRowNorms(ws, p.K.by_row, ws.scratch_n[0].get(), ws.scratch_m[0].get(), l2 != 0, ws.scratch_m[1].get());
      ws.scratch_m[1].Download(out, p.m, ws.stream);
    }
    PDLP_CUDA_CHECK(cudaGetLastError());
    ws.Sync();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_stats(pdlp_b200_device_problem* h, uint8_t** stats, size_t* stats_len) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || stats == nullptr || stats_len == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    SolveLog carrier;  // QuadraticProgramStats travels inside a SolveLog (one codec entry point)
    carrier.original_problem_stats.mut() = h->pre->ComputeStats();
    const std::string bytes = Serialize(carrier);
    *stats = DupBytes(bytes);
    *stats_len = bytes.size();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_rescale(pdlp_b200_device_problem* h, int l_inf_ruiz_iterations, int l2_norm_rescaling, double* row_scale, double* col_scale,
                                   double* scaled_values) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    PrimalDualHybridGradientParams params;
    params.l_inf_ruiz_iterations = l_inf_ruiz_iterations;
    params.l2_norm_rescaling = l2_norm_rescaling != 0;
    h->pre->RescaleForTest(params);
    if (row_scale != nullptr) p.row_scale.Download(row_scale, p.m, ws.stream);
    if (col_scale != nullptr) p.col_scale.Download(col_scale, p.n, ws.stream);
    if (scaled_values != nullptr) p.K.by_col.val.Download(scaled_values, p.K.by_col.nnz, ws.stream);
    PDLP_CUDA_CHECK(cudaGetLastError());
    ws.Sync();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_kkt_stats(pdlp_b200_device_problem* h, const uint8_t* params_bytes, size_t params_len, const double* x, const double* y,
                                     uint8_t** stats, size_t* stats_len) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || stats == nullptr || stats_len == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PrimalDualHybridGradientParams params;
    if (!ParseParams(params_bytes, params_len, &params)) return Fail(PDLP_B200_BAD_ARGUMENT, "unparsable params");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    DVec dx(p.n > 0 ? p.n : 1), dy(p.m > 0 ? p.m : 1);
    dx.Upload(x, p.n, ws.stream);
    dy.Upload(y, p.m, ws.stream);
    IterationStats out;
    ConvergenceInformation ci;
    InfeasibilityInformation ii;
    PointMetadata md;
    h->pre->ComputeConvergenceAndInfeasibility(params, dx.get(), dy.get(), dx.get(), dy.get(), POINT_TYPE_CURRENT_ITERATE, &ci, &ii, &md);
    out.convergence_information.push_back(ci);
    out.infeasibility_information.push_back(ii);
    out.point_metadata.push_back(md);
    const std::string bytes = Serialize(out);
    *stats = DupBytes(bytes);
    *stats_len = bytes.size();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_reduced_costs(pdlp_b200_device_problem* h, const uint8_t* params_bytes, size_t params_len, const double* x, const double* y,
                                         int use_zero_primal_objective, double* reduced_costs) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || reduced_costs == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PrimalDualHybridGradientParams params;
    if (!ParseParams(params_bytes, params_len, &params)) return Fail(PDLP_B200_BAD_ARGUMENT, "unparsable params");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    DVec dx(p.n > 0 ? p.n : 1), dy(p.m > 0 ? p.m : 1), rc(p.n > 0 ? p.n : 1);
    dx.Upload(x, p.n, ws.stream);
    dy.Upload(y, p.m, ws.stream);
    ComputeReducedCosts(ws, p, dx.get(), dy.get(), use_zero_primal_objective != 0, params.handle_some_primal_gradients_on_finite_bounds_as_residuals, rc.get());
    rc.Download(reduced_costs, p.n, ws.stream);
    PDLP_CUDA_CHECK(cudaGetLastError());
    ws.Sync();
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_localized_bounds(pdlp_b200_device_problem* h, const double* x, const double* y, double primal_weight, double radius,
                                            int use_diagonal, double diagonal_tol, double* out) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || out == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    DVec dx(p.n > 0 ? p.n : 1), dy(p.m > 0 ? p.m : 1);
    dx.Upload(x, p.n, ws.stream);
    dy.Upload(y, p.m, ws.stream);
    const LocalizedBounds b = ComputeLocalizedLagrangianBounds(ws, p, dx.get(), dy.get(), nullptr, primal_weight, radius, use_diagonal != 0, diagonal_tol, nullptr);
    out[0] = b.lagrangian_value; out[1] = b.lower_bound; out[2] = b.upper_bound; out[3] = b.radius;
    return PDLP_B200_OK;
  });
}

pdlp_b200_status pdlp_b200_localized_bounds_max_norm(pdlp_b200_device_problem* h, const double* x, const double* y, double primal_weight, double radius, double* out) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || out == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    DVec dx(p.n > 0 ? p.n : 1), dy(p.m > 0 ? p.m : 1);
    dx.Upload(x, p.n, ws.stream);
    dy.Upload(y, p.m, ws.stream);
    const LocalizedBounds b = ComputeLocalizedLagrangianBounds(ws, p, dx.get(), dy.get(), nullptr, primal_weight, radius, false, 0.0, nullptr, /*max_norm=*/true);
    out[0] = b.lagrangian_value; out[1] = b.lower_bound; out[2] = b.upper_bound; out[3] = b.radius;
    return PDLP_B200_OK;
  });
}

static pdlp_b200_status TrialStepsImpl(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, double* x, double* y,
                                       double* state_out, double* seconds, long long* launches, double* kernel_seconds);

pdlp_b200_status pdlp_b200_trial_steps(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, double* x, double* y,
                                       double* state_out, double* seconds, long long* launches) {
  return TrialStepsImpl(h, steps, step_size, primal_weight, x, y, state_out, seconds, launches, nullptr);
}

pdlp_b200_status pdlp_b200_trial_steps_profiled(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, const double* x,
                                                const double* y, double* kernel_seconds) {
  if (kernel_seconds == nullptr) return Fail(PDLP_B200_BAD_ARGUMENT, "kernel_seconds is null");
  std::vector<double> xs, ys;
  double* xp = nullptr;
  double* yp = nullptr;
  if (h != nullptr && x != nullptr) { xs.assign(x, x + h->pre->problem().n); xp = xs.data(); }
  if (h != nullptr && y != nullptr) { ys.assign(y, y + h->pre->problem().m); yp = ys.data(); }
  return TrialStepsImpl(h, steps, step_size, primal_weight, xp, yp, nullptr, nullptr, nullptr, kernel_seconds);
}

static pdlp_b200_status TrialStepsImpl(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, double* x, double* y,
                                       double* state_out, double* seconds, long long* launches, double* kernel_seconds) {
  return Guard([&]() -> pdlp_b200_status {
    if (h == nullptr || steps < 0) return Fail(PDLP_B200_BAD_ARGUMENT, "bad argument");
    PDLP_CUDA_CHECK(cudaSetDevice(h->device));
    Workspace& ws = h->pre->ws();
    DeviceProblem& p = h->pre->problem();
    if (!h->iterate) {
      h->iterate = std::make_unique<Iterate>();
      h->iterate->Init(p.n, p.m);
    }
    Iterate& it = *h->iterate;
    if (x != nullptr) it.x[0].Upload(x, p.n, ws.stream); else it.x[0].Zero(ws.stream);
    if (y != nullptr) it.y[0].Upload(y, p.m, ws.stream); else it.y[0].Zero(ws.stream);
    it.avg_x.Zero(ws.stream);
    it.avg_y.Zero(ws.stream);
    RowDot(ws, p.K.by_col, it.y[0].get(), it.kty[0].get());
    StepState st{};
    st.active = 1;
    st.target_iteration = 2147483647;
    st.step_size = step_size;
    st.primal_weight = primal_weight;
    st.ratio_last_two_step_sizes = 1.0;
    st.kkt_pass_limit = std::numeric_limits<double>::infinity();
    st.passes_per_rejected_step = 1.0;
    st.reduction_exponent = 0.3;
    st.growth_exponent = 0.6;
    st.mp_tau = -1.0;
    st.mp_new_tau = -1.0;
    PDLP_CUDA_CHECK(cudaMemcpyAsync(it.state.get(), &st, sizeof(st), cudaMemcpyHostToDevice, ws.stream));
    ws.Sync();
    cudaEvent_t e0, e1;
    PDLP_CUDA_CHECK(cudaEventCreate(&e0));
    PDLP_CUDA_CHECK(cudaEventCreate(&e1));
    const long long before = ws.launches;
    std::vector<cudaEvent_t> marks;
    if (kernel_seconds != nullptr) ws.profile = &marks;
    std::vector<cudaEvent_t> peer_marks;
    if (kernel_seconds != nullptr && std::getenv("PDLP_B200_DEBUG_TIMING") != nullptr) ws.profile_peer = &peer_marks;
    PDLP_CUDA_CHECK(cudaEventRecord(e0, ws.stream));
    EnqueueAdaptiveTrials(ws, p, it, steps, false);
    PDLP_CUDA_CHECK(cudaEventRecord(e1, ws.stream));
    ws.profile = nullptr;
    ws.profile_peer = nullptr;
    PDLP_CUDA_CHECK(cudaMemcpyAsync(&st, it.state.get(), sizeof(st), cudaMemcpyDeviceToHost, ws.stream));
    ws.Sync();
    if (peer_marks.size() >= 4 && peer_marks.size() % 4 == 0) {  // the fused peer pass: before, after push + private rows, after signal, after wait + finish
      double phase[3] = {0, 0, 0};
      const size_t calls = peer_marks.size() / 4;
      for (size_t c = 0; c < calls; ++c)
        for (int k = 0; k < 3; ++k) {
          float kms = 0;
          cudaEventElapsedTime(&kms, peer_marks[4 * c + k], peer_marks[4 * c + k + 1]);
          phase[k] += kms * 1e3 / calls;
        }
      std::fprintf(stderr, "[pdlp_b200] rank %d K x pass in peer mode, us per step: push + private rows %.1f, signal %.1f, wait + finish %.1f\n", ws.comm != nullptr ? ws.comm->rank() : 0,
                   phase[0], phase[1], phase[2]);
    }
    for (cudaEvent_t e : peer_marks) cudaEventDestroy(e);
    float ms = 0;
    PDLP_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (seconds != nullptr) *seconds = ms * 1e-3;
    if (launches != nullptr) *launches = ws.launches - before;
    if (kernel_seconds != nullptr) {
      // 6 marks per trial: before K1, after K1, K2, K3, K4, K5.
      for (int k = 0; k < 5; ++k) kernel_seconds[k] = 0.0;
      const int trials = static_cast<int>(marks.size() / 6);
      for (int t = 0; t < trials; ++t)
        for (int k = 0; k < 5; ++k) {
          float kms = 0;
          PDLP_CUDA_CHECK(cudaEventElapsedTime(&kms, marks[6 * t + k], marks[6 * t + k + 1]));
          kernel_seconds[k] += kms * 1e-3;
        }
      for (int k = 0; k < 5; ++k) kernel_seconds[k] /= (trials > 0 ? trials : 1);
      for (cudaEvent_t e : marks) cudaEventDestroy(e);
    }
    if (x != nullptr) it.x[st.cur].Download(x, p.n, ws.stream);
    if (y != nullptr) it.y[st.cur].Download(y, p.m, ws.stream);
    ws.Sync();
    if (state_out != nullptr) {
      state_out[0] = st.step_size; state_out[1] = st.iterations_completed; state_out[2] = st.num_rejected_steps;
      state_out[3] = st.last_movement; state_out[4] = st.last_nonlinearity;
      state_out[5] = static_cast<double>(st.total_trials); state_out[6] = st.inner; state_out[7] = st.force_numerical;
    }
    return PDLP_B200_OK;
  });
}

}  // extern "C"
