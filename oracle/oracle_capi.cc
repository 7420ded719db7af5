// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// C ABI over the CPU restatement so tests (ctypes) and bench.py's cpu_baseline leg can
// drive it.  Parameters arrive as "dotted.proto.field=value" lines (the Python side
// flattens a PrimalDualHybridGradientParams message); the solve log goes back as JSON
// with the solve_log.proto field names.
#include <cstring>
#include <map>

#include "oracle_pdhg.h"

using namespace oracle;

namespace {

std::string Num(double v) {
  if (std::isnan(v)) return "NaN";
  if (std::isinf(v)) return v > 0 ? "Infinity" : "-Infinity";
  char b[40];
  std::snprintf(b, sizeof b, "%.17g", v);
  return b;
}
std::string Str(const std::string& s) {
  std::string o = "\"";
  for (char c : s) {
    if (c == '"' || c == '\\') { o += '\\'; o += c; }
    else if (c == '\n') o += "\\n";
    else o += c;
  }
  return o + "\"";
}
struct Obj {
  std::string s = "{";
  bool first = true;
  Obj& raw(const char* k, const std::string& v) { if (!first) s += ","; first = false; s += "\""; s += k; s += "\":"; s += v; return *this; }
  Obj& num(const char* k, double v) { return raw(k, Num(v)); }
  Obj& i64(const char* k, int64_t v) { return raw(k, std::to_string(v)); }
  Obj& str(const char* k, const std::string& v) { return raw(k, Str(v)); }
  std::string done() { return s + "}"; }
};
template <typename T, typename F> std::string Arr(const std::vector<T>& v, F f) {
  std::string s = "[";
  for (size_t i = 0; i < v.size(); ++i) { if (i) s += ","; s += f(v[i]); }
  return s + "]";
}

std::string J(const ConvergenceInformation& c) {
  return Obj().i64("candidate_type", c.candidate_type).num("primal_objective", c.primal_objective).num("dual_objective", c.dual_objective)
      .num("corrected_dual_objective", c.corrected_dual_objective).num("l_inf_primal_residual", c.l_inf_primal_residual).num("l2_primal_residual", c.l2_primal_residual)
      .num("l_inf_componentwise_primal_residual", c.l_inf_componentwise_primal_residual).num("l_inf_dual_residual", c.l_inf_dual_residual)
      .num("l2_dual_residual", c.l2_dual_residual).num("l_inf_componentwise_dual_residual", c.l_inf_componentwise_dual_residual)
      .num("l_inf_primal_variable", c.l_inf_primal_variable).num("l2_primal_variable", c.l2_primal_variable).num("l_inf_dual_variable", c.l_inf_dual_variable)
      .num("l2_dual_variable", c.l2_dual_variable).done();
}
std::string J(const InfeasibilityInformation& c) {
  return Obj().num("max_primal_ray_infeasibility", c.max_primal_ray_infeasibility).num("primal_ray_linear_objective", c.primal_ray_linear_objective)
      .num("primal_ray_quadratic_norm", c.primal_ray_quadratic_norm).num("max_dual_ray_infeasibility", c.max_dual_ray_infeasibility)
      .num("dual_ray_objective", c.dual_ray_objective).i64("candidate_type", c.candidate_type).done();
}
std::string J(const PointMetadata& m) {
  Obj o;
  o.i64("point_type", m.point_type).raw("random_primal_projections", Arr(m.random_primal_projections, Num)).raw("random_dual_projections", Arr(m.random_dual_projections, Num));
  if (m.has_active)
    o.i64("active_primal_variable_count", m.active_primal_variable_count).i64("active_dual_variable_count", m.active_dual_variable_count)
        .i64("active_primal_variable_change", m.active_primal_variable_change).i64("active_dual_variable_change", m.active_dual_variable_change);
  return o.done();
}
std::string J(const IterationStats& s) {
  return Obj().i64("iteration_number", s.iteration_number)
      .raw("convergence_information", Arr(s.convergence_information, [](const ConvergenceInformation& c) { return J(c); }))
      .raw("infeasibility_information", Arr(s.infeasibility_information, [](const InfeasibilityInformation& c) { return J(c); }))
      .raw("point_metadata", Arr(s.point_metadata, [](const PointMetadata& c) { return J(c); }))
      .num("cumulative_kkt_matrix_passes", s.cumulative_kkt_matrix_passes).i64("cumulative_rejected_steps", s.cumulative_rejected_steps)
      .num("cumulative_time_sec", s.cumulative_time_sec).i64("restart_used", s.restart_used).num("step_size", s.step_size).num("primal_weight", s.primal_weight).done();
}
std::string J(const QpStats& s) {
  return Obj().i64("num_variables", s.num_variables).i64("num_constraints", s.num_constraints)
      .num("constraint_matrix_col_min_l_inf_norm", s.constraint_matrix_col_min_l_inf_norm).num("constraint_matrix_row_min_l_inf_norm", s.constraint_matrix_row_min_l_inf_norm)
      .i64("constraint_matrix_num_nonzeros", s.constraint_matrix_num_nonzeros).num("constraint_matrix_abs_max", s.constraint_matrix_abs_max)
      .num("constraint_matrix_abs_min", s.constraint_matrix_abs_min).num("constraint_matrix_abs_avg", s.constraint_matrix_abs_avg).num("constraint_matrix_l2_norm", s.constraint_matrix_l2_norm)
      .num("combined_bounds_max", s.combined_bounds_max).num("combined_bounds_min", s.combined_bounds_min).num("combined_bounds_avg", s.combined_bounds_avg)
      .num("combined_bounds_l2_norm", s.combined_bounds_l2_norm).num("combined_variable_bounds_max", s.combined_variable_bounds_max)
      .num("combined_variable_bounds_min", s.combined_variable_bounds_min).num("combined_variable_bounds_avg", s.combined_variable_bounds_avg)
      .num("combined_variable_bounds_l2_norm", s.combined_variable_bounds_l2_norm).i64("variable_bound_gaps_num_finite", s.variable_bound_gaps_num_finite)
      .num("variable_bound_gaps_max", s.variable_bound_gaps_max).num("variable_bound_gaps_min", s.variable_bound_gaps_min).num("variable_bound_gaps_avg", s.variable_bound_gaps_avg)
      .num("variable_bound_gaps_l2_norm", s.variable_bound_gaps_l2_norm).num("objective_vector_abs_max", s.objective_vector_abs_max)
      .num("objective_vector_abs_min", s.objective_vector_abs_min).num("objective_vector_abs_avg", s.objective_vector_abs_avg).num("objective_vector_l2_norm", s.objective_vector_l2_norm)
      .i64("objective_matrix_num_nonzeros", s.objective_matrix_num_nonzeros).num("objective_matrix_abs_max", s.objective_matrix_abs_max)
      .num("objective_matrix_abs_min", s.objective_matrix_abs_min).num("objective_matrix_abs_avg", s.objective_matrix_abs_avg).num("objective_matrix_l2_norm", s.objective_matrix_l2_norm).done();
}
std::string J(const FeasibilityPolishingDetails& d) {
  const DetailedCriteria& c = d.detailed_criteria;
  const std::string detailed = Obj().num("eps_optimal_primal_residual_absolute", c.eps_optimal_primal_residual_absolute)
      .num("eps_optimal_primal_residual_relative", c.eps_optimal_primal_residual_relative).num("eps_optimal_dual_residual_absolute", c.eps_optimal_dual_residual_absolute)
      .num("eps_optimal_dual_residual_relative", c.eps_optimal_dual_residual_relative).num("eps_optimal_objective_gap_absolute", c.eps_optimal_objective_gap_absolute)
      .num("eps_optimal_objective_gap_relative", c.eps_optimal_objective_gap_relative).done();
  const std::string criteria = Obj().raw("detailed_optimality_criteria", detailed).i64("iteration_limit", d.iteration_limit).done();
  return Obj().i64("polishing_phase_type", d.polishing_phase_type).i64("main_iteration_count", d.main_iteration_count)
      .raw("params", Obj().raw("termination_criteria", criteria).done()).i64("termination_reason", d.termination_reason)
      .i64("iteration_count", d.iteration_count).num("solve_time_sec", d.solve_time_sec).raw("solution_stats", J(d.solution_stats)).i64("solution_type", d.solution_type)
      .raw("iteration_stats", Arr(d.iteration_stats, [](const IterationStats& s) { return J(s); })).done();
}
std::string J(const SolveLog& l) {
  Obj o;
  o.str("instance_name", l.instance_name).i64("termination_reason", l.termination_reason).str("termination_string", l.termination_string)
      .i64("iteration_count", l.iteration_count).num("preprocessing_time_sec", l.preprocessing_time_sec).num("solve_time_sec", l.solve_time_sec)
      .raw("solution_stats", J(l.solution_stats)).i64("solution_type", l.solution_type)
      .raw("iteration_stats", Arr(l.iteration_stats, [](const IterationStats& s) { return J(s); }));
  if (l.has_original_stats) o.raw("original_problem_stats", J(l.original_problem_stats));
  if (l.has_preprocessed_stats) o.raw("preprocessed_problem_stats", J(l.preprocessed_problem_stats));
  o.raw("feasibility_polishing_details", Arr(l.feasibility_polishing_details, [](const FeasibilityPolishingDetails& d) { return J(d); }));
  return o.done();
}

char* Dup(const std::string& s) {
  char* p = static_cast<char*>(std::malloc(s.size() + 1));
  std::memcpy(p, s.c_str(), s.size() + 1);
  return p;
}

// "a.b=value" lines -> Params.  Unknown keys abort (a typo must not pass silently).
Params ParseParams(const char* text) {
  Params p;
  if (text == nullptr) return p;
  std::istringstream in(text);
  std::string line;
  TerminationCriteria& tc = p.termination_criteria;
  while (std::getline(in, line)) {
    if (line.empty()) continue;
    const size_t eq = line.find('=');
    ORACLE_CHECK(eq != std::string::npos);
    const std::string k = line.substr(0, eq), v = line.substr(eq + 1);
    auto D = [&] { return std::strtod(v.c_str(), nullptr); };
    auto I = [&] { return static_cast<int>(std::strtoll(v.c_str(), nullptr, 10)); };
    auto B = [&] { return v == "1" || v == "true" || v == "True"; };
    const std::string T = "termination_criteria.";
    const std::string S = T + "simple_optimality_criteria", Dt = T + "detailed_optimality_criteria";
    if (k == T + "optimality_norm") tc.optimality_norm = I();
    else if (k == S) tc.has_simple = true;
    else if (k == S + ".eps_optimal_absolute") { tc.has_simple = true; tc.simple_eps_abs = D(); }
    else if (k == S + ".eps_optimal_relative") { tc.has_simple = true; tc.simple_eps_rel = D(); }
    else if (k == Dt) tc.has_detailed = true;
    else if (k == Dt + ".eps_optimal_primal_residual_absolute") { tc.has_detailed = true; tc.detailed.eps_optimal_primal_residual_absolute = D(); }
    else if (k == Dt + ".eps_optimal_primal_residual_relative") { tc.has_detailed = true; tc.detailed.eps_optimal_primal_residual_relative = D(); }
    else if (k == Dt + ".eps_optimal_dual_residual_absolute") { tc.has_detailed = true; tc.detailed.eps_optimal_dual_residual_absolute = D(); }
    else if (k == Dt + ".eps_optimal_dual_residual_relative") { tc.has_detailed = true; tc.detailed.eps_optimal_dual_residual_relative = D(); }
    else if (k == Dt + ".eps_optimal_objective_gap_absolute") { tc.has_detailed = true; tc.detailed.eps_optimal_objective_gap_absolute = D(); }
    else if (k == Dt + ".eps_optimal_objective_gap_relative") { tc.has_detailed = true; tc.detailed.eps_optimal_objective_gap_relative = D(); }
    else if (k == T + "eps_optimal_absolute") { tc.has_eps_optimal_absolute = true; tc.eps_optimal_absolute = D(); }
    else if (k == T + "eps_optimal_relative") { tc.has_eps_optimal_relative = true; tc.eps_optimal_relative = D(); }
    else if (k == T + "eps_primal_infeasible") tc.eps_primal_infeasible = D();
    else if (k == T + "eps_dual_infeasible") tc.eps_dual_infeasible = D();
    else if (k == T + "time_sec_limit") tc.time_sec_limit = D();
    else if (k == T + "iteration_limit") tc.iteration_limit = I();
    else if (k == T + "kkt_matrix_pass_limit") tc.kkt_matrix_pass_limit = D();
    else if (k == "termination_criteria") {}
    else if (k == "num_threads") p.num_threads = I();
    else if (k == "num_shards") p.num_shards = I();
    else if (k == "record_iteration_stats") p.record_iteration_stats = B();
    else if (k == "verbosity_level") p.verbosity_level = I();
    else if (k == "log_interval_seconds") p.log_interval_seconds = D();
    else if (k == "major_iteration_frequency") p.major_iteration_frequency = I();
    else if (k == "termination_check_frequency") p.termination_check_frequency = I();
    else if (k == "restart_strategy") p.restart_strategy = I();
    else if (k == "primal_weight_update_smoothing") p.primal_weight_update_smoothing = D();
    else if (k == "initial_primal_weight") { p.has_initial_primal_weight = true; p.initial_primal_weight = D(); }
    else if (k == "presolve_options" ) {}
    else if (k == "presolve_options.use_glop") p.use_glop = B();
    else if (k == "l_inf_ruiz_iterations") p.l_inf_ruiz_iterations = I();
    else if (k == "l2_norm_rescaling") p.l2_norm_rescaling = B();
    else if (k == "sufficient_reduction_for_restart") p.sufficient_reduction_for_restart = D();
    else if (k == "necessary_reduction_for_restart") p.necessary_reduction_for_restart = D();
    else if (k == "linesearch_rule") p.linesearch_rule = I();
    else if (k == "adaptive_linesearch_parameters" || k == "malitsky_pock_parameters") {}
    else if (k == "adaptive_linesearch_parameters.step_size_reduction_exponent") p.step_size_reduction_exponent = D();
    else if (k == "adaptive_linesearch_parameters.step_size_growth_exponent") p.step_size_growth_exponent = D();
    else if (k == "malitsky_pock_parameters.step_size_downscaling_factor") p.step_size_downscaling_factor = D();
    else if (k == "malitsky_pock_parameters.linesearch_contraction_factor") p.linesearch_contraction_factor = D();
    else if (k == "malitsky_pock_parameters.step_size_interpolation") p.step_size_interpolation = D();
    else if (k == "initial_step_size_scaling") p.initial_step_size_scaling = D();
    else if (k == "random_projection_seeds") { std::istringstream ss(v); std::string t; while (std::getline(ss, t, ',')) if (!t.empty()) p.random_projection_seeds.push_back(std::atoi(t.c_str())); }
    else if (k == "infinite_constraint_bound_threshold") p.infinite_constraint_bound_threshold = D();
    else if (k == "handle_some_primal_gradients_on_finite_bounds_as_residuals") p.handle_some_primal_gradients_on_finite_bounds_as_residuals = B();
    else if (k == "use_diagonal_qp_trust_region_solver") p.use_diagonal_qp_trust_region_solver = B();
    else if (k == "diagonal_qp_trust_region_solver_tolerance") p.diagonal_qp_trust_region_solver_tolerance = D();
    else if (k == "use_feasibility_polishing") p.use_feasibility_polishing = B();
    else { std::fprintf(stderr, "oracle: unknown parameter key '%s'\n", k.c_str()); std::abort(); }
  }
  return p;
}

Qp MakeQp(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val, const double* c, const double* q,
          const double* vlb, const double* vub, const double* clb, const double* cub, double offset, double scale) {
  Qp qp;
  qp.K.rows = m; qp.K.cols = n;
  qp.K.ptr.assign(col_ptr, col_ptr + n + 1);
  qp.K.idx.assign(row_idx, row_idx + nnz);
  qp.K.val.assign(val, val + nnz);
  qp.c.assign(c, c + n);
  if (q != nullptr) qp.q = Vec(q, q + n);
  qp.vlb.assign(vlb, vlb + n); qp.vub.assign(vub, vub + n);
  qp.clb.assign(clb, clb + m); qp.cub.assign(cub, cub + m);
  qp.objective_offset = offset; qp.objective_scaling_factor = scale;
  return qp;
}

}  // namespace

extern "C" {

typedef void (*oracle_callback)(int iteration_type, const char* iteration_stats_json, void* user);

void oracle_free(void* p) { std::free(p); }

// Full solve.  x_out/y_out/rc_out may be shorter than n/m on error results; *sizes
// receive the lengths actually written (0 for error results).
int oracle_solve(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val, const double* c, const double* q_diag,
                 const double* vlb, const double* vub, const double* clb, const double* cub, double objective_offset, double objective_scaling_factor,
                 const char* params_text, const double* x0, const double* y0, const volatile int* interrupt, oracle_callback cb, void* cb_user, double* x_out,
                 double* y_out, double* rc_out, int64_t* sizes /*3*/, char** log_json) {
  Qp qp = MakeQp(n, m, nnz, col_ptr, row_idx, val, c, q_diag, vlb, vub, clb, cub, objective_offset, objective_scaling_factor);
  const Params params = ParseParams(params_text);
  Vec vx0, vy0;
  if (x0 != nullptr && y0 != nullptr) { vx0.assign(x0, x0 + n); vy0.assign(y0, y0 + m); }
  Callback callback;
  if (cb != nullptr) callback = [&](const IterationCallbackInfo& info) { const std::string s = J(*info.iteration_stats); cb(static_cast<int>(info.iteration_type), s.c_str(), cb_user); };
  SolverResult r = PrimalDualHybridGradient(std::move(qp), params, x0 ? &vx0 : nullptr, y0 ? &vy0 : nullptr, interrupt, callback);
  sizes[0] = r.primal_solution.size(); sizes[1] = r.dual_solution.size(); sizes[2] = r.reduced_costs.size();
  std::copy(r.primal_solution.begin(), r.primal_solution.end(), x_out);
  std::copy(r.dual_solution.begin(), r.dual_solution.end(), y_out);
  std::copy(r.reduced_costs.begin(), r.reduced_costs.end(), rc_out);
  *log_json = Dup(J(r.solve_log));
  return 0;
}

// Bench support: seconds of `steps` iterations of the step loop alone after `warmup` untimed ones
// (no termination checks, no restarts; oracle_pdhg_impl.h:TimeSteps).  out = {seconds, accepted steps,
// rejected steps, preprocessing seconds}.  Returns 0, or 1 when preprocessing rejects the problem or a
// step forces numerical termination.
int oracle_time_steps(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val, const double* c, const double* q_diag,
                      const double* vlb, const double* vub, const double* clb, const double* cub, double objective_offset, double objective_scaling_factor,
                      const char* params_text, int warmup, int steps, double* out /*4*/) {
  Qp qp = MakeQp(n, m, nnz, col_ptr, row_idx, val, c, q_diag, vlb, vub, clb, cub, objective_offset, objective_scaling_factor);
  const Params params = ParseParams(params_text);
  PreprocessSolver solver(std::move(qp), params);
  StepTiming t;
  const bool ok = solver.TimeSteps(params, warmup, steps, t);
  out[0] = t.seconds; out[1] = t.accepted_steps; out[2] = t.rejected_steps; out[3] = t.preprocessing_seconds;
  return ok && t.completed ? 0 : 1;
}

// ---- unit entry points for the golden tests ----
void oracle_validate_params(const char* params_text, char** message) { *message = Dup(ValidateParams(ParseParams(params_text))); }

// Termination predicates (termination.cc:161-219,239-271) on flat records.
//   conv: n_conv rows of {candidate_type, primal_objective, dual_objective, l_inf_primal_residual, l2_primal_residual,
//         l_inf_componentwise_primal_residual, l_inf_dual_residual, l2_dual_residual, l_inf_componentwise_dual_residual}
//   inf:  n_inf rows of {candidate_type, max_primal_ray_infeasibility, primal_ray_linear_objective, primal_ray_quadratic_norm,
//         max_dual_ray_infeasibility, dual_ray_objective}
//   work: {iteration_number, cumulative_kkt_matrix_passes, cumulative_time_sec};  out: {reason, point_type} (0, 0 = nullopt)
void oracle_check_termination(const char* params_text, int which, int force_numerical, const volatile int* interrupt, const double* norms, int n_conv,
                              const double* conv, int n_inf, const double* inf, const double* work, int* out, double* relative_residuals) {
  const Params params = ParseParams(params_text);
  IterationStats s;
  for (int i = 0; i < n_conv; ++i) {
    const double* r = conv + 9 * i;
    ConvergenceInformation c;
    c.candidate_type = static_cast<int>(r[0]); c.primal_objective = r[1]; c.dual_objective = r[2];
    c.l_inf_primal_residual = r[3]; c.l2_primal_residual = r[4]; c.l_inf_componentwise_primal_residual = r[5];
    c.l_inf_dual_residual = r[6]; c.l2_dual_residual = r[7]; c.l_inf_componentwise_dual_residual = r[8];
    s.convergence_information.push_back(c);
  }
  for (int i = 0; i < n_inf; ++i) {
    const double* r = inf + 6 * i;
    InfeasibilityInformation c;
    c.candidate_type = static_cast<int>(r[0]); c.max_primal_ray_infeasibility = r[1]; c.primal_ray_linear_objective = r[2];
    c.primal_ray_quadratic_norm = r[3]; c.max_dual_ray_infeasibility = r[4]; c.dual_ray_objective = r[5];
    s.infeasibility_information.push_back(c);
  }
  s.iteration_number = static_cast<int>(work[0]); s.cumulative_kkt_matrix_passes = work[1]; s.cumulative_time_sec = work[2];
  const BoundNorms bn{norms[0], norms[1], norms[2], norms[3]};
  std::optional<TerminationReasonAndPointType> r;
  if (which == 0) r = CheckIterateTerminationCriteria(params.termination_criteria, s, bn, force_numerical != 0);
  else r = CheckSimpleTerminationCriteria(params.termination_criteria, s, interrupt);
  out[0] = r.has_value() ? r->reason : 0;
  out[1] = r.has_value() ? r->type : 0;
  if (relative_residuals != nullptr && n_conv > 0) {
    const RelativeConvergenceInformation rel = ComputeRelativeResiduals(EffectiveOptimalityCriteria(params.termination_criteria), s.convergence_information[0], bn);
    relative_residuals[0] = rel.relative_l_inf_primal_residual; relative_residuals[1] = rel.relative_l2_primal_residual;
    relative_residuals[2] = rel.relative_l_inf_dual_residual; relative_residuals[3] = rel.relative_l2_dual_residual;
    relative_residuals[4] = rel.relative_optimality_gap;
  }
}

int oracle_shard_starts(int64_t n, int num_shards, const int64_t* masses /*nullable: uniform*/, int64_t* starts_out /*<= num_shards+1*/) {
  std::unique_ptr<Sharder> sh;
  if (masses) sh = std::make_unique<Sharder>(n, num_shards, nullptr, [&](int64_t i) { return masses[i]; });
  else sh = std::make_unique<Sharder>(n, num_shards, nullptr);
  for (size_t i = 0; i < sh->Starts().size(); ++i) starts_out[i] = sh->Starts()[i];
  return sh->NumShards();
}

// The sharded BLAS-1 of sharder.h:241-335 by name, on vectors of length n with `num_shards` shards, for the small
// examples of sharder_test.cc:260-405.  Vector results go to out_vec (n doubles), scalar results are returned.
double oracle_blas1(const char* op, int64_t n, int num_shards, const double* a, const double* b, double scalar, double* out_vec) {
  Sharder sh(n, num_shards, nullptr);
  const std::string name(op);
  const Vec va = a != nullptr ? Vec(a, a + n) : Vec(), vb = b != nullptr ? Vec(b, b + n) : Vec();
  auto store = [&](const Vec& v) { std::copy(v.begin(), v.end(), out_vec); return 0.0; };
  if (name == "SetZero") { Vec d = va; SetZero(sh, d); return store(d); }
  if (name == "ZeroVector") return store(ZeroVector(sh));
  if (name == "OnesVector") return store(OnesVector(sh));
  if (name == "AddScaledVector") { Vec d = va; AddScaledVector(scalar, vb, sh, d); return store(d); }
  if (name == "AssignVector") { Vec d; AssignVector(va, sh, d); return store(d); }
  if (name == "CloneVector") return store(CloneVector(va, sh));
  if (name == "CoefficientWiseProductInPlace") { Vec d = va; CwiseProductInPlace(vb, sh, d); return store(d); }
  if (name == "CoefficientWiseQuotientInPlace") { Vec d = va; CwiseQuotientInPlace(vb, sh, d); return store(d); }
  if (name == "Dot") return Dot(va, vb, sh);
  if (name == "LInfNorm") return LInfNorm(va, sh);
  if (name == "L1Norm") return L1Norm(va, sh);
  if (name == "SquaredNorm") return SquaredNorm(va, sh);
  if (name == "Norm") return Norm(va, sh);
  if (name == "SquaredDistance") return SquaredDistance(va, vb, sh);
  if (name == "Distance") return Distance(va, vb, sh);
  if (name == "ScaledLInfNorm") return ScaledLInfNorm(va, vb, sh);
  if (name == "ScaledSquaredNorm") return ScaledSquaredNorm(va, vb, sh);
  if (name == "ScaledNorm") return ScaledNorm(va, vb, sh);
  ORACLE_CHECK(false && "unknown BLAS-1 op");
  return 0.0;
}

// ShardedWeightedAverage (sharded_optimization_utils.cc:43-79) fed `count` vectors of length n with their weights;
// clear_after_add >= 0: Clear() after that many adds.  meta_out = {HasNonzeroWeight, NumTerms}.
void oracle_weighted_average(int64_t n, int num_shards, int count, const double* vectors, const double* weights, int clear_after_add, double* avg_out,
                             double* meta_out) {
  Sharder sh(n, num_shards, nullptr);
  WeightedAverage avg(&sh);
  for (int k = 0; k < count; ++k) {
    avg.Add(Vec(vectors + k * n, vectors + (k + 1) * n), weights[k]);
    if (k + 1 == clear_after_add) avg.Clear();
  }
  const Vec out = avg.ComputeAverage();
  std::copy(out.begin(), out.end(), avg_out);
  meta_out[0] = avg.HasNonzeroWeight() ? 1.0 : 0.0;
  meta_out[1] = avg.NumTerms();
}

struct OracleProblem { ShardedQp* sq; Vec col_scale, row_scale; };

void* oracle_problem_create(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val, const double* c, const double* q_diag,
                            const double* vlb, const double* vub, const double* clb, const double* cub, double offset, double scale, int num_threads, int num_shards) {
  auto* p = new OracleProblem;
  p->sq = new ShardedQp(MakeQp(n, m, nnz, col_ptr, row_idx, val, c, q_diag, vlb, vub, clb, cub, offset, scale), num_threads, num_shards);
  p->col_scale = OnesVector(p->sq->PrimalSharder());
  p->row_scale = OnesVector(p->sq->DualSharder());
  return p;
}
void oracle_problem_destroy(void* h) { auto* p = static_cast<OracleProblem*>(h); delete p->sq; delete p; }

// which: 0 -> K^T y (out length n), 1 -> K x (out length m).
void oracle_matvec(void* h, int which, const double* v, double* out) {
  auto* p = static_cast<OracleProblem*>(h);
  const ShardedQp& sq = *p->sq;
  if (which == 0) { Vec y(v, v + sq.DualSize()); Vec r = TransposedMatVec(sq.qp().K, y, sq.KSharder()); std::copy(r.begin(), r.end(), out); }
  else { Vec x(v, v + sq.PrimalSize()); Vec r = TransposedMatVec(sq.Kt(), x, sq.KtSharder()); std::copy(r.begin(), r.end(), out); }
}
// Column norms of diag(row_scale) K diag(col_scale); l2 = 0 -> Linf, 1 -> L2.
void oracle_col_norms(void* h, int l2, const double* row_scale, const double* col_scale, double* out) {
  auto* p = static_cast<OracleProblem*>(h);
  const ShardedQp& sq = *p->sq;
  Vec r(row_scale, row_scale + sq.DualSize()), c(col_scale, col_scale + sq.PrimalSize());
  Vec o = l2 ? ScaledColL2Norm(sq.qp().K, r, c, sq.KSharder()) : ScaledColLInfNorm(sq.qp().K, r, c, sq.KSharder());
  std::copy(o.begin(), o.end(), out);
}
void oracle_stats(void* h, char** json) { *json = Dup(J(ComputeStats(*static_cast<OracleProblem*>(h)->sq))); }

// Rescales the problem in place; scaling vectors accumulate into the handle and are copied out.
void oracle_rescale(void* h, int l_inf_ruiz_iterations, int l2_norm_rescaling, double* row_scale_out, double* col_scale_out) {
  auto* p = static_cast<OracleProblem*>(h);
  ScalingVectors s = ApplyRescaling(l_inf_ruiz_iterations, l2_norm_rescaling != 0, *p->sq);
  for (size_t i = 0; i < s.row.size(); ++i) p->row_scale[i] *= s.row[i];
  for (size_t i = 0; i < s.col.size(); ++i) p->col_scale[i] *= s.col[i];
  std::copy(s.row.begin(), s.row.end(), row_scale_out);
  std::copy(s.col.begin(), s.col.end(), col_scale_out);
}
// One scaling pass without applying it: which = 0 -> Linf Ruiz (iters), 1 -> L2.
void oracle_scaling_vectors(void* h, int which, int iters, double* row_scale_inout, double* col_scale_inout) {
  auto* p = static_cast<OracleProblem*>(h);
  const ShardedQp& sq = *p->sq;
  Vec r(row_scale_inout, row_scale_inout + sq.DualSize()), c(col_scale_inout, col_scale_inout + sq.PrimalSize());
  if (which == 0) LInfRuizRescaling(sq, iters, r, c); else L2NormRescaling(sq, r, c);
  std::copy(r.begin(), r.end(), row_scale_inout);
  std::copy(c.begin(), c.end(), col_scale_inout);
}
// ShardedQuadraticProgram::RescaleQuadraticProgram with given vectors (sharded_quadratic_program.cc:134-167) and
// ReplaceLargeConstraintBoundsWithInfinity (:119-130), for the goldens of sharded_quadratic_program_test.cc.
void oracle_rescale_with(void* h, const double* col_scale, const double* row_scale) {
  auto* p = static_cast<OracleProblem*>(h);
  const Vec col(col_scale, col_scale + p->sq->PrimalSize()), row(row_scale, row_scale + p->sq->DualSize());
  p->sq->Rescale(col, row);
  for (size_t i = 0; i < row.size(); ++i) p->row_scale[i] *= row[i];
  for (size_t i = 0; i < col.size(); ++i) p->col_scale[i] *= col[i];
}
void oracle_replace_large_constraint_bounds(void* h, double threshold) { static_cast<OracleProblem*>(h)->sq->ReplaceLargeConstraintBoundsWithInfinity(threshold); }
// Values of the transposed copy, in its own compressed order (rows of K, columns ascending).
void oracle_problem_get_transposed_values(void* h, double* val) {
  const Csc& kt = static_cast<OracleProblem*>(h)->sq->Kt();
  std::copy(kt.val.begin(), kt.val.end(), val);
}
// SetRandomProjections (iteration_stats.cc:598-610): one primal and one dual projection per seed.
void oracle_random_projections(void* h, const double* x, const double* y, const int* seeds, int count, double* primal_out, double* dual_out) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  const Vec vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  for (int k = 0; k < count; ++k) {
    std::mt19937 gen(seeds[k]);
    primal_out[k] = RandomProjection(vx, sq.PrimalSharder(), gen);
    dual_out[k] = RandomProjection(vy, sq.DualSharder(), gen);
  }
}
// ProjectToPrimalVariableBounds / ProjectToDualVariableBounds (sharded_optimization_utils.cc:701-750), in place.
void oracle_project(void* h, double* x /*nullable*/, double* y /*nullable*/) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  if (x != nullptr) {
    Vec v(x, x + sq.PrimalSize());
    ProjectToPrimalVariableBounds(sq, v);
    std::copy(v.begin(), v.end(), x);
  }
  if (y != nullptr) {
    Vec v(y, y + sq.DualSize());
    ProjectToDualVariableBounds(sq, v);
    std::copy(v.begin(), v.end(), y);
  }
}
// Reads the (possibly rescaled) problem data back: any pointer may be null.
void oracle_problem_get(void* h, double* val, double* c, double* q, double* vlb, double* vub, double* clb, double* cub) {
  const Qp& qp = static_cast<OracleProblem*>(h)->sq->qp();
  if (val) std::copy(qp.K.val.begin(), qp.K.val.end(), val);
  if (c) std::copy(qp.c.begin(), qp.c.end(), c);
  if (q && qp.q) std::copy(qp.q->begin(), qp.q->end(), q);
  if (vlb) std::copy(qp.vlb.begin(), qp.vlb.end(), vlb);
  if (vub) std::copy(qp.vub.begin(), qp.vub.end(), vub);
  if (clb) std::copy(qp.clb.begin(), qp.clb.end(), clb);
  if (cub) std::copy(qp.cub.begin(), qp.cub.end(), cub);
}

void oracle_convergence_info(void* h, const char* params_text, const double* col_scale, const double* row_scale, const double* x, const double* y,
                             double cw_primal_offset, double cw_dual_offset, int candidate_type, char** json) {
  auto* p = static_cast<OracleProblem*>(h);
  const ShardedQp& sq = *p->sq;
  const Params params = ParseParams(params_text);
  Vec cs(col_scale, col_scale + sq.PrimalSize()), rs(row_scale, row_scale + sq.DualSize()), vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  *json = Dup(J(ComputeConvergenceInformation(params, sq, cs, rs, vx, vy, cw_primal_offset, cw_dual_offset, candidate_type)));
}
void oracle_infeasibility_info(void* h, const char* params_text, const double* col_scale, const double* row_scale, const double* x, const double* y, int candidate_type,
                               char** json) {
  auto* p = static_cast<OracleProblem*>(h);
  const ShardedQp& sq = *p->sq;
  const Params params = ParseParams(params_text);
  Vec cs(col_scale, col_scale + sq.PrimalSize()), rs(row_scale, row_scale + sq.DualSize()), vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  *json = Dup(J(ComputeInfeasibilityInformation(params, sq, cs, rs, vx, vy, candidate_type)));
}
void oracle_reduced_costs(void* h, const char* params_text, const double* x, const double* y, int zero_objective, double* rc_out) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  const Params params = ParseParams(params_text);
  Vec vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  Vec rc = ReducedCosts(params, sq, vx, vy, zero_objective != 0);
  std::copy(rc.begin(), rc.end(), rc_out);
}
// Lagrangian gradients: primal (needs K^T y) and dual (needs K x); returns the value parts.
void oracle_lagrangian_gradients(void* h, const double* x, const double* y, double* primal_grad, double* dual_grad, double* values /*2*/) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  Vec vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  const Vec kty = TransposedMatVec(sq.qp().K, vy, sq.KSharder());
  const Vec kx = TransposedMatVec(sq.Kt(), vx, sq.KtSharder());
  const LagrangianPart pp = ComputePrimalGradient(sq, vx, kty), dp = ComputeDualGradient(sq, vy, kx);
  std::copy(pp.gradient.begin(), pp.gradient.end(), primal_grad);
  std::copy(dp.gradient.begin(), dp.gradient.end(), dual_grad);
  values[0] = pp.value; values[1] = dp.value;
}
// out = {lagrangian_value, lower_bound, upper_bound, radius}.  norm: 0 max, 1 Euclidean.
void oracle_localized_bounds(void* h, const double* x, const double* y, int norm, double primal_weight, double radius, int use_diagonal_qp_solver, double tol, double* out) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  Vec vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize());
  const LocalizedLagrangianBounds b = ComputeLocalizedLagrangianBounds(sq, vx, vy, norm == 0 ? PrimalDualNorm::kMaxNorm : PrimalDualNorm::kEuclideanNorm, primal_weight, radius,
                                                                       nullptr, nullptr, use_diagonal_qp_solver != 0, tol);
  out[0] = b.lagrangian_value; out[1] = b.lower_bound; out[2] = b.upper_bound; out[3] = b.radius;
}
// The same with the caller's K x and K^T y (trust_region.h: primal_product / dual_product, used by the solver's restart test).
void oracle_localized_bounds_cached(void* h, const double* x, const double* y, const double* kx, const double* kty, int norm, double primal_weight, double radius,
                                    int use_diagonal_qp_solver, double tol, double* out) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  const Vec vx(x, x + sq.PrimalSize()), vy(y, y + sq.DualSize()), vkx(kx, kx + sq.DualSize()), vkty(kty, kty + sq.PrimalSize());
  const LocalizedLagrangianBounds b = ComputeLocalizedLagrangianBounds(sq, vx, vy, norm == 0 ? PrimalDualNorm::kMaxNorm : PrimalDualNorm::kEuclideanNorm, primal_weight, radius,
                                                                       &vkx, &vkty, use_diagonal_qp_solver != 0, tol);
  out[0] = b.lagrangian_value; out[1] = b.lower_bound; out[2] = b.upper_bound; out[3] = b.radius;
}
// Plain vector trust-region problem.  diag nullable (linear objective).  out2 = {step_size, objective_value}.
void oracle_trust_region(int64_t n, const double* obj, const double* diag, const double* lb, const double* ub, const double* center, const double* w, double radius,
                         int num_shards, double tol, double* solution, double* out2) {
  Sharder sh(n, num_shards, nullptr);
  Vec o(obj, obj + n), l(lb, lb + n), u(ub, ub + n), c(center, center + n), ww(w, w + n);
  TrustRegionResult r;
  if (diag) { Vec d(diag, diag + n); r = SolveDiagonalTrustRegion(o, d, l, u, c, ww, radius, sh, tol); }
  else r = SolveTrustRegion(o, l, u, c, ww, radius, sh);
  std::copy(r.solution.begin(), r.solution.end(), solution);
  out2[0] = r.solution_step_size; out2[1] = r.objective_value;
}
double oracle_sigma_max(void* h, double rel_err, double fail_prob, int seed, int* iterations) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  std::mt19937 gen(seed);
  const auto r = EstimateMaximumSingularValue(sq, nullptr, nullptr, rel_err, fail_prob, gen);
  if (iterations) *iterations = r.num_iterations;
  return r.singular_value;
}

// The same with the active subspaces of a primal and / or a dual point (either may be NULL).
double oracle_sigma_max_at(void* h, const double* x, const double* y, double rel_err, double fail_prob, int seed, int* iterations) {
  const ShardedQp& sq = *static_cast<OracleProblem*>(h)->sq;
  std::mt19937 gen(seed);
  Vec vx, vy;
  if (x) vx.assign(x, x + sq.PrimalSize());
  if (y) vy.assign(y, y + sq.DualSize());
  const auto r = EstimateMaximumSingularValueOfConstraintMatrix(sq, x ? &vx : nullptr, y ? &vy : nullptr, rel_err, fail_prob, gen);
  if (iterations) *iterations = r.num_iterations;
  return r.singular_value;
}

}  // extern "C"
