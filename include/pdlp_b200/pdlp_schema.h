// Plain-struct mirrors of the two PDLP protobuf schemas plus a small proto2 wire codec,
// so parameter / solve-log bytes are interchangeable with the reference's
//   ortools/pdlp/solvers.proto:20-481   (PrimalDualHybridGradientParams, TerminationCriteria, ...)
//   ortools/pdlp/solve_log.proto:24-450 (SolveLog, IterationStats, ConvergenceInformation, ...)
// Same field names, numbers, types and defaults.  libprotobuf is not linked: each message
// lists its fields once in Visit(); a writer, a reader and a schema describer walk that list.
// The Python package builds its message classes from Describe(), so there is one source of
// truth for the schema.
#ifndef PDLP_B200_SCHEMA_H_
#define PDLP_B200_SCHEMA_H_

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

namespace pdlp_b200 {

// ---- enums (numbers as in the .proto files) ----
enum OptimalityNorm { OPTIMALITY_NORM_UNSPECIFIED = 0, OPTIMALITY_NORM_L_INF = 1, OPTIMALITY_NORM_L2 = 2, OPTIMALITY_NORM_L_INF_COMPONENTWISE = 3 };
enum RestartStrategy { RESTART_STRATEGY_UNSPECIFIED = 0, NO_RESTARTS = 1, EVERY_MAJOR_ITERATION = 2, ADAPTIVE_HEURISTIC = 3, ADAPTIVE_DISTANCE_BASED = 4 };
enum LinesearchRule { LINESEARCH_RULE_UNSPECIFIED = 0, ADAPTIVE_LINESEARCH_RULE = 1, MALITSKY_POCK_LINESEARCH_RULE = 2, CONSTANT_STEP_SIZE_RULE = 3 };
enum RestartChoice { RESTART_CHOICE_UNSPECIFIED = 0, RESTART_CHOICE_NO_RESTART = 1, RESTART_CHOICE_WEIGHTED_AVERAGE_RESET = 2, RESTART_CHOICE_RESTART_TO_AVERAGE = 3 };
enum PointType { POINT_TYPE_UNSPECIFIED = 0, POINT_TYPE_CURRENT_ITERATE = 1, POINT_TYPE_ITERATE_DIFFERENCE = 2, POINT_TYPE_AVERAGE_ITERATE = 3, POINT_TYPE_NONE = 4,
                 POINT_TYPE_PRESOLVER_SOLUTION = 5, POINT_TYPE_FEASIBILITY_POLISHING_SOLUTION = 6 };
enum TerminationReason { TERMINATION_REASON_UNSPECIFIED = 0, TERMINATION_REASON_OPTIMAL = 1, TERMINATION_REASON_PRIMAL_INFEASIBLE = 2, TERMINATION_REASON_DUAL_INFEASIBLE = 3,
                         TERMINATION_REASON_TIME_LIMIT = 4, TERMINATION_REASON_ITERATION_LIMIT = 5, TERMINATION_REASON_NUMERICAL_ERROR = 6, TERMINATION_REASON_OTHER = 7,
                         TERMINATION_REASON_KKT_MATRIX_PASS_LIMIT = 8, TERMINATION_REASON_INVALID_PROBLEM = 9, TERMINATION_REASON_INVALID_PARAMETER = 10,
                         TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE = 11, TERMINATION_REASON_INTERRUPTED_BY_USER = 12, TERMINATION_REASON_INVALID_INITIAL_SOLUTION = 13 };
enum PolishingPhaseType { POLISHING_PHASE_TYPE_UNSPECIFIED = 0, POLISHING_PHASE_TYPE_PRIMAL_FEASIBILITY = 1, POLISHING_PHASE_TYPE_DUAL_FEASIBILITY = 2 };

const char* TerminationReasonName(int r);
const char* PointTypeName(int t);

// ---- field wrappers ----
template <typename T> struct Opt {  // optional scalar with a schema default
  T v; bool has = false;
  Opt(T def = T()) : v(def) {}
  operator T() const { return v; }
  T get() const { return v; }
  void set(T x) { v = x; has = true; }
  Opt& operator=(T x) { set(x); return *this; }
};
template <typename E> struct OptEnum {  // optional enum (varint on the wire)
  int v; bool has = false;
  OptEnum(E def = E()) : v(static_cast<int>(def)) {}
  operator int() const { return v; }
  void set(int x) { v = x; has = true; }
  OptEnum& operator=(int x) { set(x); return *this; }
  static const char* TypeName();
};
template <typename M> struct OptMsg {  // optional sub-message
  M v; bool has = false;
  M* operator->() { return &v; }
  const M* operator->() const { return &v; }
  M& mut() { has = true; return v; }
  const M& get() const { return v; }
};
template <typename M> using Rep = std::vector<M>;  // repeated sub-message
template <typename T> struct Packed { std::vector<T> v; };  // repeated packed scalar
struct RawMsg { std::string bytes; bool has = false; const char* type_name; explicit RawMsg(const char* t) : type_name(t) {} };

constexpr double kSchemaInf = std::numeric_limits<double>::infinity();

// ---- solvers.proto ----
struct SimpleOptimalityCriteria {
  static constexpr const char* kName = "TerminationCriteria.SimpleOptimalityCriteria";
  Opt<double> eps_optimal_absolute{1.0e-6}, eps_optimal_relative{1.0e-6};
  template <class V> void Visit(V& v) { v.f(1, "eps_optimal_absolute", eps_optimal_absolute); v.f(2, "eps_optimal_relative", eps_optimal_relative); }
};
struct DetailedOptimalityCriteria {
  static constexpr const char* kName = "TerminationCriteria.DetailedOptimalityCriteria";
  Opt<double> eps_optimal_primal_residual_absolute{1.0e-6}, eps_optimal_primal_residual_relative{1.0e-6};
  Opt<double> eps_optimal_dual_residual_absolute{1.0e-6}, eps_optimal_dual_residual_relative{1.0e-6};
  Opt<double> eps_optimal_objective_gap_absolute{1.0e-6}, eps_optimal_objective_gap_relative{1.0e-6};
  template <class V> void Visit(V& v) {
    v.f(1, "eps_optimal_primal_residual_absolute", eps_optimal_primal_residual_absolute);
    v.f(2, "eps_optimal_primal_residual_relative", eps_optimal_primal_residual_relative);
    v.f(3, "eps_optimal_dual_residual_absolute", eps_optimal_dual_residual_absolute);
    v.f(4, "eps_optimal_dual_residual_relative", eps_optimal_dual_residual_relative);
    v.f(5, "eps_optimal_objective_gap_absolute", eps_optimal_objective_gap_absolute);
    v.f(6, "eps_optimal_objective_gap_relative", eps_optimal_objective_gap_relative);
  }
};
struct TerminationCriteria {
  static constexpr const char* kName = "TerminationCriteria";
  OptEnum<OptimalityNorm> optimality_norm{OPTIMALITY_NORM_L2};
  OptMsg<SimpleOptimalityCriteria> simple_optimality_criteria;      // oneof optimality_criteria
  OptMsg<DetailedOptimalityCriteria> detailed_optimality_criteria;  // oneof optimality_criteria
  Opt<double> eps_optimal_absolute{1.0e-6}, eps_optimal_relative{1.0e-6};  // deprecated pair
  Opt<double> eps_primal_infeasible{1.0e-8}, eps_dual_infeasible{1.0e-8};
  Opt<double> time_sec_limit{kSchemaInf};
  Opt<int32_t> iteration_limit{2147483647};
  Opt<double> kkt_matrix_pass_limit{kSchemaInf};
  template <class V> void Visit(V& v) {
    v.f(1, "optimality_norm", optimality_norm);
    v.oneof_begin("optimality_criteria");
    v.f(9, "simple_optimality_criteria", simple_optimality_criteria);
    v.f(10, "detailed_optimality_criteria", detailed_optimality_criteria);
    v.oneof_end();
    v.f(2, "eps_optimal_absolute", eps_optimal_absolute);
    v.f(3, "eps_optimal_relative", eps_optimal_relative);
    v.f(4, "eps_primal_infeasible", eps_primal_infeasible);
    v.f(5, "eps_dual_infeasible", eps_dual_infeasible);
    v.f(6, "time_sec_limit", time_sec_limit);
    v.f(7, "iteration_limit", iteration_limit);
    v.f(8, "kkt_matrix_pass_limit", kkt_matrix_pass_limit);
  }
  // Setting one member of the oneof clears the other (proto2 semantics).
  void OneofFixup(int just_set) { if (just_set == 9) detailed_optimality_criteria.has = false; if (just_set == 10) simple_optimality_criteria.has = false; }
};
struct AdaptiveLinesearchParams {
  static constexpr const char* kName = "AdaptiveLinesearchParams";
  Opt<double> step_size_reduction_exponent{0.3}, step_size_growth_exponent{0.6};
  template <class V> void Visit(V& v) { v.f(1, "step_size_reduction_exponent", step_size_reduction_exponent); v.f(2, "step_size_growth_exponent", step_size_growth_exponent); }
};
struct MalitskyPockParams {
  static constexpr const char* kName = "MalitskyPockParams";
  Opt<double> step_size_downscaling_factor{0.7}, linesearch_contraction_factor{0.99}, step_size_interpolation{1.0};
  template <class V> void Visit(V& v) {
    v.f(1, "step_size_downscaling_factor", step_size_downscaling_factor);
    v.f(2, "linesearch_contraction_factor", linesearch_contraction_factor);
    v.f(3, "step_size_interpolation", step_size_interpolation);
  }
};
struct PresolveOptions {
  static constexpr const char* kName = "PrimalDualHybridGradientParams.PresolveOptions";
  Opt<bool> use_glop{false};
  RawMsg glop_parameters{".operations_research.glop.GlopParameters"};  // carried opaquely
  template <class V> void Visit(V& v) { v.f(1, "use_glop", use_glop); v.f(2, "glop_parameters", glop_parameters); }
};
struct PrimalDualHybridGradientParams {
  static constexpr const char* kName = "PrimalDualHybridGradientParams";
  OptMsg<TerminationCriteria> termination_criteria;
  Opt<int32_t> num_threads{1}, num_shards{0};
  Opt<bool> record_iteration_stats{false};
  Opt<int32_t> verbosity_level{0};
  Opt<double> log_interval_seconds{0.0};
  Opt<int32_t> major_iteration_frequency{64}, termination_check_frequency{64};
  OptEnum<RestartStrategy> restart_strategy{ADAPTIVE_HEURISTIC};
  Opt<double> primal_weight_update_smoothing{0.5};
  Opt<double> initial_primal_weight{0.0};
  OptMsg<PresolveOptions> presolve_options;
  Opt<int32_t> l_inf_ruiz_iterations{5};
  Opt<bool> l2_norm_rescaling{true};
  Opt<double> sufficient_reduction_for_restart{0.1}, necessary_reduction_for_restart{0.9};
  OptEnum<LinesearchRule> linesearch_rule{ADAPTIVE_LINESEARCH_RULE};
  OptMsg<AdaptiveLinesearchParams> adaptive_linesearch_parameters;
  OptMsg<MalitskyPockParams> malitsky_pock_parameters;
  Opt<double> initial_step_size_scaling{1.0};
  Packed<int32_t> random_projection_seeds;
  Opt<double> infinite_constraint_bound_threshold{kSchemaInf};
  Opt<bool> handle_some_primal_gradients_on_finite_bounds_as_residuals{true};
  Opt<bool> use_diagonal_qp_trust_region_solver{false};
  Opt<double> diagonal_qp_trust_region_solver_tolerance{1.0e-8};
  Opt<bool> use_feasibility_polishing{false};
  template <class V> void Visit(V& v) {
    v.f(1, "termination_criteria", termination_criteria);
    v.f(2, "num_threads", num_threads);
    v.f(27, "num_shards", num_shards);
    v.f(3, "record_iteration_stats", record_iteration_stats);
    v.f(26, "verbosity_level", verbosity_level);
    v.f(31, "log_interval_seconds", log_interval_seconds);
    v.f(4, "major_iteration_frequency", major_iteration_frequency);
    v.f(5, "termination_check_frequency", termination_check_frequency);
    v.f(6, "restart_strategy", restart_strategy);
    v.f(7, "primal_weight_update_smoothing", primal_weight_update_smoothing);
    v.f(8, "initial_primal_weight", initial_primal_weight);
    v.f(16, "presolve_options", presolve_options);
    v.f(9, "l_inf_ruiz_iterations", l_inf_ruiz_iterations);
    v.f(10, "l2_norm_rescaling", l2_norm_rescaling);
    v.f(11, "sufficient_reduction_for_restart", sufficient_reduction_for_restart);
    v.f(17, "necessary_reduction_for_restart", necessary_reduction_for_restart);
    v.f(12, "linesearch_rule", linesearch_rule);
    v.f(18, "adaptive_linesearch_parameters", adaptive_linesearch_parameters);
    v.f(19, "malitsky_pock_parameters", malitsky_pock_parameters);
    v.f(25, "initial_step_size_scaling", initial_step_size_scaling);
    v.f(28, "random_projection_seeds", random_projection_seeds);
    v.f(22, "infinite_constraint_bound_threshold", infinite_constraint_bound_threshold);
    v.f(29, "handle_some_primal_gradients_on_finite_bounds_as_residuals", handle_some_primal_gradients_on_finite_bounds_as_residuals);
    v.f(23, "use_diagonal_qp_trust_region_solver", use_diagonal_qp_trust_region_solver);
    v.f(24, "diagonal_qp_trust_region_solver_tolerance", diagonal_qp_trust_region_solver_tolerance);
    v.f(30, "use_feasibility_polishing", use_feasibility_polishing);
  }
};

// ---- solve_log.proto ----
struct QuadraticProgramStats {
  static constexpr const char* kName = "QuadraticProgramStats";
  Opt<int64_t> num_variables, num_constraints;
  Opt<double> constraint_matrix_col_min_l_inf_norm, constraint_matrix_row_min_l_inf_norm;
  Opt<int64_t> constraint_matrix_num_nonzeros;
  Opt<double> constraint_matrix_abs_max, constraint_matrix_abs_min, constraint_matrix_abs_avg, constraint_matrix_l2_norm;
  Opt<double> combined_bounds_max, combined_bounds_min, combined_bounds_avg, combined_bounds_l2_norm;
  Opt<double> combined_variable_bounds_max, combined_variable_bounds_min, combined_variable_bounds_avg, combined_variable_bounds_l2_norm;
  Opt<int64_t> variable_bound_gaps_num_finite;
  Opt<double> variable_bound_gaps_max, variable_bound_gaps_min, variable_bound_gaps_avg, variable_bound_gaps_l2_norm;
  Opt<double> objective_vector_abs_max, objective_vector_abs_min, objective_vector_abs_avg, objective_vector_l2_norm;
  Opt<int64_t> objective_matrix_num_nonzeros;
  Opt<double> objective_matrix_abs_max, objective_matrix_abs_min, objective_matrix_abs_avg, objective_matrix_l2_norm;
  template <class V> void Visit(V& v) {
    v.f(1, "num_variables", num_variables); v.f(2, "num_constraints", num_constraints);
    v.f(3, "constraint_matrix_col_min_l_inf_norm", constraint_matrix_col_min_l_inf_norm);
    v.f(4, "constraint_matrix_row_min_l_inf_norm", constraint_matrix_row_min_l_inf_norm);
    v.f(5, "constraint_matrix_num_nonzeros", constraint_matrix_num_nonzeros);
    v.f(6, "constraint_matrix_abs_max", constraint_matrix_abs_max); v.f(7, "constraint_matrix_abs_min", constraint_matrix_abs_min);
    v.f(8, "constraint_matrix_abs_avg", constraint_matrix_abs_avg); v.f(25, "constraint_matrix_l2_norm", constraint_matrix_l2_norm);
    v.f(9, "combined_bounds_max", combined_bounds_max); v.f(10, "combined_bounds_min", combined_bounds_min);
    v.f(11, "combined_bounds_avg", combined_bounds_avg); v.f(24, "combined_bounds_l2_norm", combined_bounds_l2_norm);
    v.f(28, "combined_variable_bounds_max", combined_variable_bounds_max); v.f(29, "combined_variable_bounds_min", combined_variable_bounds_min);
    v.f(30, "combined_variable_bounds_avg", combined_variable_bounds_avg); v.f(31, "combined_variable_bounds_l2_norm", combined_variable_bounds_l2_norm);
    v.f(12, "variable_bound_gaps_num_finite", variable_bound_gaps_num_finite);
    v.f(13, "variable_bound_gaps_max", variable_bound_gaps_max); v.f(14, "variable_bound_gaps_min", variable_bound_gaps_min);
    v.f(15, "variable_bound_gaps_avg", variable_bound_gaps_avg); v.f(26, "variable_bound_gaps_l2_norm", variable_bound_gaps_l2_norm);
    v.f(16, "objective_vector_abs_max", objective_vector_abs_max); v.f(17, "objective_vector_abs_min", objective_vector_abs_min);
    v.f(18, "objective_vector_abs_avg", objective_vector_abs_avg); v.f(23, "objective_vector_l2_norm", objective_vector_l2_norm);
    v.f(19, "objective_matrix_num_nonzeros", objective_matrix_num_nonzeros);
    v.f(20, "objective_matrix_abs_max", objective_matrix_abs_max); v.f(21, "objective_matrix_abs_min", objective_matrix_abs_min);
    v.f(22, "objective_matrix_abs_avg", objective_matrix_abs_avg); v.f(27, "objective_matrix_l2_norm", objective_matrix_l2_norm);
  }
};
struct ConvergenceInformation {
  static constexpr const char* kName = "ConvergenceInformation";
  OptEnum<PointType> candidate_type;
  Opt<double> primal_objective, dual_objective, corrected_dual_objective;
  Opt<double> l_inf_primal_residual, l2_primal_residual, l_inf_componentwise_primal_residual;
  Opt<double> l_inf_dual_residual, l2_dual_residual, l_inf_componentwise_dual_residual;
  Opt<double> l_inf_primal_variable, l2_primal_variable, l_inf_dual_variable, l2_dual_variable;
  template <class V> void Visit(V& v) {
    v.f(1, "candidate_type", candidate_type);
    v.f(2, "primal_objective", primal_objective); v.f(3, "dual_objective", dual_objective); v.f(4, "corrected_dual_objective", corrected_dual_objective);
    v.f(5, "l_inf_primal_residual", l_inf_primal_residual); v.f(6, "l2_primal_residual", l2_primal_residual);
    v.f(24, "l_inf_componentwise_primal_residual", l_inf_componentwise_primal_residual);
    v.f(7, "l_inf_dual_residual", l_inf_dual_residual); v.f(8, "l2_dual_residual", l2_dual_residual);
    v.f(25, "l_inf_componentwise_dual_residual", l_inf_componentwise_dual_residual);
    v.f(14, "l_inf_primal_variable", l_inf_primal_variable); v.f(15, "l2_primal_variable", l2_primal_variable);
    v.f(16, "l_inf_dual_variable", l_inf_dual_variable); v.f(17, "l2_dual_variable", l2_dual_variable);
  }
};
struct InfeasibilityInformation {
  static constexpr const char* kName = "InfeasibilityInformation";
  Opt<double> max_primal_ray_infeasibility, primal_ray_linear_objective, primal_ray_quadratic_norm, max_dual_ray_infeasibility, dual_ray_objective;
  OptEnum<PointType> candidate_type;
  template <class V> void Visit(V& v) {
    v.f(1, "max_primal_ray_infeasibility", max_primal_ray_infeasibility); v.f(2, "primal_ray_linear_objective", primal_ray_linear_objective);
    v.f(3, "primal_ray_quadratic_norm", primal_ray_quadratic_norm); v.f(4, "max_dual_ray_infeasibility", max_dual_ray_infeasibility);
    v.f(5, "dual_ray_objective", dual_ray_objective); v.f(6, "candidate_type", candidate_type);
  }
};
struct PointMetadata {
  static constexpr const char* kName = "PointMetadata";
  OptEnum<PointType> point_type;
  Packed<double> random_primal_projections, random_dual_projections;
  Opt<int64_t> active_primal_variable_count, active_dual_variable_count, active_primal_variable_change, active_dual_variable_change;
  template <class V> void Visit(V& v) {
    v.f(1, "point_type", point_type);
    v.f(2, "random_primal_projections", random_primal_projections); v.f(3, "random_dual_projections", random_dual_projections);
    v.f(4, "active_primal_variable_count", active_primal_variable_count); v.f(5, "active_dual_variable_count", active_dual_variable_count);
    v.f(6, "active_primal_variable_change", active_primal_variable_change); v.f(7, "active_dual_variable_change", active_dual_variable_change);
  }
};
struct IterationStats {
  static constexpr const char* kName = "IterationStats";
  Opt<int32_t> iteration_number;
  Rep<ConvergenceInformation> convergence_information;
  Rep<InfeasibilityInformation> infeasibility_information;
  Rep<PointMetadata> point_metadata;
  Opt<double> cumulative_kkt_matrix_passes;
  Opt<int32_t> cumulative_rejected_steps;
  Opt<double> cumulative_time_sec;
  OptEnum<RestartChoice> restart_used;
  Opt<double> step_size, primal_weight;
  template <class V> void Visit(V& v) {
    v.f(1, "iteration_number", iteration_number);
    v.f(2, "convergence_information", convergence_information);
    v.f(3, "infeasibility_information", infeasibility_information);
    v.f(11, "point_metadata", point_metadata);
    v.f(4, "cumulative_kkt_matrix_passes", cumulative_kkt_matrix_passes);
    v.f(5, "cumulative_rejected_steps", cumulative_rejected_steps);
    v.f(6, "cumulative_time_sec", cumulative_time_sec);
    v.f(7, "restart_used", restart_used);
    v.f(8, "step_size", step_size);
    v.f(9, "primal_weight", primal_weight);
  }
};
struct FeasibilityPolishingDetails {
  static constexpr const char* kName = "FeasibilityPolishingDetails";
  OptEnum<PolishingPhaseType> polishing_phase_type;
  Opt<int32_t> main_iteration_count;
  OptMsg<PrimalDualHybridGradientParams> params;
  OptEnum<TerminationReason> termination_reason;
  Opt<int32_t> iteration_count;
  Opt<double> solve_time_sec;
  OptMsg<IterationStats> solution_stats;
  OptEnum<PointType> solution_type;
  Rep<IterationStats> iteration_stats;
  template <class V> void Visit(V& v) {
    v.f(1, "polishing_phase_type", polishing_phase_type); v.f(2, "main_iteration_count", main_iteration_count);
    v.f(3, "params", params); v.f(4, "termination_reason", termination_reason); v.f(5, "iteration_count", iteration_count);
    v.f(6, "solve_time_sec", solve_time_sec); v.f(7, "solution_stats", solution_stats); v.f(8, "solution_type", solution_type);
    v.f(9, "iteration_stats", iteration_stats);
  }
};
struct SolveLog {
  static constexpr const char* kName = "SolveLog";
  Opt<std::string> instance_name;
  OptMsg<PrimalDualHybridGradientParams> params;
  OptEnum<TerminationReason> termination_reason;
  Opt<std::string> termination_string;
  Opt<int32_t> iteration_count;
  Opt<double> preprocessing_time_sec, solve_time_sec;
  OptMsg<IterationStats> solution_stats;
  OptEnum<PointType> solution_type;
  Rep<IterationStats> iteration_stats;
  OptMsg<QuadraticProgramStats> original_problem_stats, preprocessed_problem_stats;
  Rep<FeasibilityPolishingDetails> feasibility_polishing_details;
  template <class V> void Visit(V& v) {
    v.f(1, "instance_name", instance_name); v.f(14, "params", params); v.f(3, "termination_reason", termination_reason);
    v.f(4, "termination_string", termination_string); v.f(5, "iteration_count", iteration_count);
    v.f(13, "preprocessing_time_sec", preprocessing_time_sec); v.f(6, "solve_time_sec", solve_time_sec);
    v.f(8, "solution_stats", solution_stats); v.f(10, "solution_type", solution_type); v.f(7, "iteration_stats", iteration_stats);
    v.f(11, "original_problem_stats", original_problem_stats); v.f(12, "preprocessed_problem_stats", preprocessed_problem_stats);
    v.f(15, "feasibility_polishing_details", feasibility_polishing_details);
  }
};

// ---- wire codec entry points (pdlp_schema.cc) ----
std::string Serialize(const PrimalDualHybridGradientParams& m);
std::string Serialize(const SolveLog& m);
std::string Serialize(const IterationStats& m);
// Returns false on malformed input.
bool Parse(const void* data, size_t len, PrimalDualHybridGradientParams* out);
bool Parse(const void* data, size_t len, SolveLog* out);
bool Parse(const void* data, size_t len, IterationStats* out);
// Line-oriented schema description consumed by the Python package (see _schema.py).
std::string DescribeSchema();

}  // namespace pdlp_b200
#endif  // PDLP_B200_SCHEMA_H_
