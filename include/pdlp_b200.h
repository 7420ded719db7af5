/* C ABI of the B200-native PDLP solver: the drop-in boundary for
 *   SolverResult PrimalDualHybridGradient(QuadraticProgram, PrimalDualHybridGradientParams,
 *                                         [initial_solution], interrupt_solve, callback)
 * of /root/reference/ortools/pdlp/primal_dual_hybrid_gradient.h:140-156.
 *
 * The reference has no FFI layer of its own for this path (it is a C++ free function
 * plus two protobuf schemas; its Python binding is pybind11 over that function,
 * ortools/pdlp/python/pdlp.cc:143-150).  These entry points are what such a binding
 * would call: plain pointers and sizes, serialized protobuf messages for the two
 * schema types (wire-compatible with ortools/pdlp/solvers.proto and solve_log.proto).
 * INTEGRATION.md shows the reference-side stub.
 *
 * Error convention: solver-level failures (invalid problem / parameters / initial
 * solution) are a NORMAL return with termination_reason set in the SolveLog, exactly
 * as the reference does (primal_dual_hybrid_gradient.cc:717-725).  The return status
 * is non-zero only for ABI-level failures (null pointers, CUDA errors, out of memory).
 * There is no CPU fallback: without a usable CUDA device every solve returns
 * PDLP_B200_CUDA_ERROR.
 */
#ifndef PDLP_B200_H_
#define PDLP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  PDLP_B200_OK = 0,
  PDLP_B200_BAD_ARGUMENT = 1, /* null pointer / negative size / unparsable params bytes */
  PDLP_B200_CUDA_ERROR = 2,   /* no device, launch failure, out of device memory */
  PDLP_B200_INTERNAL_ERROR = 3
} pdlp_b200_status;

/* The QuadraticProgram of ortools/pdlp/quadratic_program.h:53-143, by borrowed
 * pointers.  The constraint matrix is compressed sparse column (the layout of a
 * compressed Eigen::SparseMatrix<double, ColMajor, int64_t>: outerIndexPtr /
 * innerIndexPtr / valuePtr).  Infinite bounds are +-INFINITY.  Nothing is modified. */
typedef struct {
  int64_t num_variables;    /* n */
  int64_t num_constraints;  /* m */
  int64_t num_nonzeros;
  const int64_t* col_ptr;   /* n + 1 */
  const int64_t* row_idx;   /* num_nonzeros */
  const double* values;     /* num_nonzeros */
  const double* objective_vector;           /* n */
  const double* objective_matrix_diagonal;  /* n, or NULL for a linear program */
  const double* variable_lower_bounds;      /* n */
  const double* variable_upper_bounds;      /* n */
  const double* constraint_lower_bounds;    /* m */
  const double* constraint_upper_bounds;    /* m */
  double objective_offset;
  double objective_scaling_factor;          /* -1 encodes maximisation */
  const char* problem_name;                 /* NULL if unnamed */
  /* Optional 32-bit form of the offsets / indices (what scipy and most CSC producers hold below 2^31
   * nonzeros): when col_ptr32 is non-NULL the pair replaces col_ptr / row_idx, which may then be NULL. */
  const int32_t* col_ptr32;                 /* n + 1 */
  const int32_t* row_idx32;                 /* num_nonzeros */
} pdlp_b200_problem;

/* IterationCallbackInfo (primal_dual_hybrid_gradient.h:62-93): iteration_type is the
 * IterationType enum value in the reference's order (:70-83: 0 kNormal, 1 kPrimalFeasibility,
 * 2 kDualFeasibility, 3 kPresolveTermination, 4 kNormalTermination,
 * 5 kFeasibilityPolishingTermination); iteration_stats is a serialized IterationStats message;
 * bound_norms = {l2_norm_primal_linear_objective, l2_norm_constraint_bounds,
 * l_inf_norm_primal_linear_objective, l_inf_norm_constraint_bounds}.  Runs synchronously
 * on the calling thread at every termination check and once at exit; it may set the
 * interrupt flag. */
typedef void (*pdlp_b200_callback)(int iteration_type, const uint8_t* iteration_stats, size_t iteration_stats_len,
                                   const double* bound_norms /*4*/, void* user);

typedef struct {
  long long kernel_launches;
  long long trial_steps;        /* accepted + rejected PDHG steps */
  double setup_seconds;         /* upload + transpose + stats + rescaling */
  double loop_seconds;          /* host wall time of the iteration loop */
  double device_loop_seconds;   /* CUDA-event time of the iteration loop */
  long long collectives;        /* exchange steps enqueued: NCCL all-reduces + peer-memory exchanges (0 on one GPU) */
  double upload_seconds;        /* host->device copies, CSR build, communicator set-up */
  int symmetric_buffers;        /* exchange buffers NCCL registered as symmetric windows (0: plain all-reduce) */
  long long peer_exchanges;     /* exchange steps that went through peer memory instead of ncclAllReduce (counted in `collectives` too) */
  int interrupts_seen_on_device; /* times the step-decision kernel itself saw the interrupt flag and stopped the loop (one GPU) */
} pdlp_b200_counters;

/* Solves `problem`.
 *   params / params_len   serialized PrimalDualHybridGradientParams (NULL / 0 = defaults)
 *   initial_primal/dual   both NULL (start at zero) or both given (sizes n and m)
 *   interrupt             NULL, or a flag tested after every iteration like the reference's interrupt_solve
 *                         (termination.cc:178): the step-decision kernel reads a host-mapped copy that the waiting
 *                         host keeps up to date, so the solve stops within one iteration of the flag being raised
 *                         (sharded solves: at the next host event, where the ranks agree on it)
 *   device                CUDA device ordinal, or -1 for the current device
 *   primal/dual/reduced   caller-allocated outputs of sizes n, m, n; written only when
 *                         the result carries a solution (*has_solution = 1)
 *   solve_log             callee-allocated serialized SolveLog; free with pdlp_b200_free
 *   counters              optional (NULL allowed)
 */
pdlp_b200_status pdlp_b200_solve(const pdlp_b200_problem* problem, const uint8_t* params, size_t params_len,
                                 const double* initial_primal_solution, const double* initial_dual_solution,
                                 const volatile int* interrupt, pdlp_b200_callback callback, void* callback_user, int device,
                                 double* primal_solution, double* dual_solution, double* reduced_costs, int* has_solution,
                                 uint8_t** solve_log, size_t* solve_log_len, pdlp_b200_counters* counters);

/* ---- column-sharded solve over several GPUs (one process per GPU) ----
 * The reference shards every pass over K by columns across CPU threads (ShardedQuadraticProgram,
 * sharded_quadratic_program.h:38-60; Sharder mass rule sharder.cc:94-138).  Here rank r of `world`
 * owns the contiguous column block [column_offset, column_offset + block.num_variables) of the
 * global problem: `block` holds those columns of K (row indices stay global, num_constraints = m)
 * and the matching slices of the objective vector / diagonal and the variable bounds; the
 * constraint bounds are given in full on every rank.  The data path has one exchange step per
 * trial step (an NCCL all-reduce of the partial K x, m doubles) plus scalar all-reduces.
 * nccl_unique_id: PDLP_B200_UNIQUE_ID_BYTES bytes from pdlp_b200_unique_id() on ONE rank,
 * distributed to the others by the caller (any channel: MPI, torch.distributed, a file).
 * world == 1 or shard == NULL: a plain single-GPU solve.
 *
 * Private constraints.  A constraint whose row of K has nonzeros in ONE rank's columns only needs
 * no exchange at all.  The caller may therefore give each rank just the constraints its columns
 * touch, renumbered locally: first the num_shared_constraints constraints that several ranks touch
 * (the same ones, in the same order, on every rank), then the rank's private ones;
 * block.num_constraints = shared + private, row indices of `block` are these local numbers, and the
 * constraint bounds / dual vectors of the call have that length and order.  Only the shared rows'
 * sums are all-reduced (8 * num_shared_constraints bytes per trial step).  A constraint no rank
 * touches may be given to any one rank as a private one.  num_shared_constraints < 0: every rank
 * holds all constraints (all of them shared). */
#define PDLP_B200_UNIQUE_ID_BYTES 128
typedef struct {
  int rank, world;
  int64_t global_num_variables;
  int64_t column_offset;
  const uint8_t* nccl_unique_id;
  int64_t num_shared_constraints;
} pdlp_b200_shard;

pdlp_b200_status pdlp_b200_unique_id(uint8_t* out /*PDLP_B200_UNIQUE_ID_BYTES*/);

/* The communicator behind a unique id is created by the first sharded call that presents the id
 * (collective: all ranks must make that call) and then kept for the life of the process: later
 * solves with the same id / rank / world reuse it.  This destroys every kept communicator (collective;
 * no solve may be running); *released (optional) = how many. */
pdlp_b200_status pdlp_b200_release_communicators(int* released);

/* Same contract as pdlp_b200_solve; every rank must call it with the same params and the same
 * presence of initial solution / interrupt / callback.  initial_primal_solution, primal_solution
 * and reduced_costs are this rank's block (block.num_variables entries); the dual vectors have
 * block.num_constraints entries (shared part identical on all ranks); the SolveLog is the whole
 * problem's and identical on all ranks (up to wall-clock fields).  The
 * interrupt flag may be raised on any rank: all ranks stop at the same iteration. */
pdlp_b200_status pdlp_b200_solve_sharded(const pdlp_b200_problem* block, const pdlp_b200_shard* shard, const uint8_t* params, size_t params_len,
                                         const double* initial_primal_solution, const double* initial_dual_solution,
                                         const volatile int* interrupt, pdlp_b200_callback callback, void* callback_user, int device,
                                         double* primal_solution, double* dual_solution, double* reduced_costs, int* has_solution,
                                         uint8_t** solve_log, size_t* solve_log_len, pdlp_b200_counters* counters);

/* ---- presolve hand-off ----
 * The reference runs Glop's preprocessor inside the solve when presolve_options.use_glop is set
 * (ApplyPresolveIfEnabled, primal_dual_hybrid_gradient.cc:1145-1202): PDHG iterates on the presolved LP while
 * convergence / infeasibility statistics, termination tests and the returned solution are those of the ORIGINAL
 * problem at the point RecoverOriginalSolution (:1288-1341) maps back, and reduced costs are recomputed on the
 * original (:1571-1585).  Glop is host code outside this library, so the presolver is the caller's:
 *   original          the problem as given (sizes of primal / dual / reduced outputs, of the statistics)
 *   presolved         what PDHG iterates on; NULL when the presolver solved the problem by itself - then
 *                     presolve_status is its verdict as a TerminationReason (GlopStatusToTerminationReason,
 *                     :1123-1143) and the result is built as in :986-1034 (solution_type
 *                     POINT_TYPE_PRESOLVER_SOLUTION, callback type kPresolveTermination; OPTIMAL only if the
 *                     recovered point really meets the optimality criteria, else NUMERICAL_ERROR)
 *   postsolve         maps an UNSCALED solution of the presolved problem (both vectors, zero-length when
 *                     presolved is NULL) to a solution of the original; called on the solving thread at every
 *                     termination check and once at the end; the library then projects onto the original's
 *                     variable bounds and dual sign constraints, as the reference does
 *   initial solution  ignored with the reference's warning (:1164-1168)
 * presolve_options.use_glop in params is not required and not consulted.  Single GPU only.  Everything
 * else as pdlp_b200_solve. */
typedef void (*pdlp_b200_postsolve)(const double* presolved_primal, const double* presolved_dual, double* original_primal, double* original_dual, void* user);
pdlp_b200_status pdlp_b200_solve_presolved(const pdlp_b200_problem* original, const pdlp_b200_problem* presolved, int presolve_status, pdlp_b200_postsolve postsolve,
                                           void* postsolve_user, const uint8_t* params, size_t params_len, const double* initial_primal_solution,
                                           const double* initial_dual_solution, const volatile int* interrupt, pdlp_b200_callback callback, void* callback_user,
                                           int device, double* primal_solution, double* dual_solution, double* reduced_costs, int* has_solution, uint8_t** solve_log,
                                           size_t* solve_log_len, pdlp_b200_counters* counters);

/* ValidatePrimalDualHybridGradientParams (solvers_proto_validation.cc:171-298): returns
 * PDLP_B200_OK and *message = NULL when valid, else *message = malloc'ed status text. */
pdlp_b200_status pdlp_b200_validate_params(const uint8_t* params, size_t params_len, char** message);

/* The termination predicates (termination.cc:161-219,239-271), host code, no device needed.
 *   which = 0: CheckIterateTerminationCriteria(params.termination_criteria, stats, bound_norms,
 *              force_numerical_termination); which = 1: CheckSimpleTerminationCriteria(criteria,
 *              stats, interrupt).
 *   stats: serialized IterationStats; bound_norms as in pdlp_b200_callback.
 *   *reason / *point_type: TerminationReason / PointType numbers, both 0 (UNSPECIFIED) when the
 *   reference would return std::nullopt.
 *   relative_residuals (optional, 5): ComputeRelativeResiduals of stats.convergence_information(0):
 *   {l_inf primal, l2 primal, l_inf dual, l2 dual, optimality gap}. */
pdlp_b200_status pdlp_b200_check_termination(const uint8_t* params, size_t params_len, const uint8_t* stats, size_t stats_len,
                                             const double* bound_norms /*4*/, int which, int force_numerical_termination,
                                             const volatile int* interrupt, int* reason, int* point_type, double* relative_residuals);

/* Host only: the iteration log table of verbosity_level >= 2 (LogIterationStatsHeader / LogIterationStats,
 * primal_dual_hybrid_gradient.cc:193-287) as text, for one serialized IterationStats: what the solve writes
 * to stderr at every termination check.  params supplies verbosity_level, use_feasibility_polishing and the
 * termination criteria; iteration_type / preferred_candidate are IterationType / PointType numbers;
 * bound_norms as in pdlp_b200_callback.  *header and *line are malloc'ed (pdlp_b200_free). */
pdlp_b200_status pdlp_b200_format_iteration_stats(const uint8_t* params, size_t params_len, const uint8_t* stats, size_t stats_len, const double* bound_norms /*4*/,
                                                  int iteration_type, int preferred_candidate, char** header, char** line);

/* Serialized PrimalDualHybridGradientParams with every field at its schema default: under proto2 presence rules that
 * is the message with no field set, so *params_len is 0 (and *params a valid pointer to free). */
pdlp_b200_status pdlp_b200_default_params(uint8_t** params, size_t* params_len);

/* Line-oriented description of the two schemas (message / field / number / type /
 * default), from which language bindings build their message classes. */
pdlp_b200_status pdlp_b200_describe_schema(char** text);

/* Text of the last ABI-level error on this thread (valid until the next call). */
const char* pdlp_b200_last_error(void);

void pdlp_b200_free(void* p);

/* ---- kernel-level entry points (unit parity tests, ncu captures) ----
 * A handle owns the device copy of a problem (both matrix orientations). */
typedef struct pdlp_b200_device_problem pdlp_b200_device_problem;

pdlp_b200_status pdlp_b200_problem_create(const pdlp_b200_problem* problem, int device, pdlp_b200_device_problem** out);
/* The handle of one rank's column block: every kernel-level entry point below then works on the
 * block (primal-side arguments have block.num_variables entries) and is collective. */
pdlp_b200_status pdlp_b200_problem_create_sharded(const pdlp_b200_problem* block, const pdlp_b200_shard* shard, int device,
                                                  pdlp_b200_device_problem** out);
void pdlp_b200_problem_destroy(pdlp_b200_device_problem* h);

/* TransposedMatrixVectorProduct (sharder.cc:163-176): which = 0 -> out[n] = K^T v[m],
 * which = 1 -> out[m] = K v[n].  Host in / host out; `repeat` launches are timed on
 * the device, *seconds_per_launch (optional) receives the CUDA-event average. */
pdlp_b200_status pdlp_b200_matvec(pdlp_b200_device_problem* h, int which, const double* v, double* out, int repeat, double* seconds_per_launch);

/* ScaledColLInfNorm / ScaledColL2Norm (sharder.cc:291-335) of diag(row_scale) K
 * diag(col_scale): which = 0 -> column norms (n), 1 -> row norms (m); l2 = 0 / 1. */
pdlp_b200_status pdlp_b200_scaled_norms(pdlp_b200_device_problem* h, int which, int l2, const double* row_scale, const double* col_scale, double* out);

/* ComputeStats (sharded_optimization_utils.cc:270-343): serialized QuadraticProgramStats. */
pdlp_b200_status pdlp_b200_stats(pdlp_b200_device_problem* h, uint8_t** stats, size_t* stats_len);

/* ApplyRescaling (sharded_optimization_utils.cc:423-444) in place on the handle;
 * outputs the cumulative scaling vectors (row: m, col: n) and, optionally, the scaled
 * matrix values in CSC order (nnz). */
pdlp_b200_status pdlp_b200_rescale(pdlp_b200_device_problem* h, int l_inf_ruiz_iterations, int l2_norm_rescaling, double* row_scale, double* col_scale,
                                   double* scaled_values);

/* ComputeConvergenceInformation + ComputeInfeasibilityInformation
 * (iteration_stats.cc:349-518) at (x, y) of the handle's current (possibly rescaled)
 * problem: serialized IterationStats with one entry of each. */
pdlp_b200_status pdlp_b200_kkt_stats(pdlp_b200_device_problem* h, const uint8_t* params, size_t params_len, const double* x, const double* y,
                                     uint8_t** stats, size_t* stats_len);

/* ReducedCosts (iteration_stats.cc:533-565). */
pdlp_b200_status pdlp_b200_reduced_costs(pdlp_b200_device_problem* h, const uint8_t* params, size_t params_len, const double* x, const double* y,
                                         int use_zero_primal_objective, double* reduced_costs);

/* ComputeLocalizedLagrangianBounds, Euclidean norm (trust_region.cc:886-1016):
 * out = {lagrangian_value, lower_bound, upper_bound, radius}. */
pdlp_b200_status pdlp_b200_localized_bounds(pdlp_b200_device_problem* h, const double* x, const double* y, double primal_weight, double radius,
                                            int use_diagonal_qp_trust_region_solver, double diagonal_qp_trust_region_solver_tolerance, double* out);

/* ComputeLocalizedLagrangianBounds with PrimalDualNorm::kMaxNorm (trust_region.cc:797-884, dispatcher :1003-1007):
 * the primal and the dual trust-region problem solved separately with radii sqrt(2) r / sqrt(w) and
 * sqrt(2) r sqrt(w).  out as above.  (PDHG itself only asks for the Euclidean norm.) */
pdlp_b200_status pdlp_b200_localized_bounds_max_norm(pdlp_b200_device_problem* h, const double* x, const double* y, double primal_weight, double radius, double* out);

/* `steps` trial steps of the adaptive rule from (x, y) with the given step size and
 * primal weight, nothing else (no restarts, no checks): the timed unit of the
 * benchmark.  Outputs the iterate after the last accepted step and the step state:
 * state_out = {step_size, iterations_completed, num_rejected_steps, last_movement,
 * last_nonlinearity, trials executed, rejected trials of the unfinished iteration (counted in
 * num_rejected_steps only once that iteration accepts a step), 1 if a trial forced numerical
 * termination and turned the rest of the batch into no-ops}; *seconds (optional) = CUDA-event time
 * of the whole batch. */
pdlp_b200_status pdlp_b200_trial_steps(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, double* x /*n, in-out*/,
                                       double* y /*m, in-out*/, double* state_out /*8*/, double* seconds, long long* launches);

/* Same as pdlp_b200_trial_steps with a CUDA event recorded around every launch:
 * kernel_seconds = average duration of {primal_step, dual_step (K x), kty_step (K^T y),
 * decide, average}.  For the per-kernel roofline table of bench.py. */
pdlp_b200_status pdlp_b200_trial_steps_profiled(pdlp_b200_device_problem* h, int steps, double step_size, double primal_weight, const double* x,
                                                const double* y, double* kernel_seconds /*5*/);

/* K v WITHOUT ANY ALLOCATION: out[rows] = K v for a CSR matrix the caller holds in device memory.
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernel as pdlp_b200_matvec(which = 1) (rowstream_kernel<RowDotOp>); the work-block table
 * that pdlp_b200_problem_create keeps in cudaMalloc'ed arrays is built on the device into a
 * file-scope static __device__ array of sizeof(int) * 2 * (PDLP_B200_NO_ALLOC_MAX_ROWS + 1) + 64 bytes
 * (about 2 MB), which every GPU that loads the library holds until the process exits
 * (or-tools_b200/csrc/pdlp_noalloc.cu).  The call runs no cudaMalloc / cudaFree / allocating library
 * routine; it enqueues on `stream` (a cudaStream_t, NULL = the default stream) and returns after the
 * result is complete.
 *   rows, cols        shape of K; rows > PDLP_B200_NO_ALLOC_MAX_ROWS is rejected with
 *                     PDLP_B200_BAD_ARGUMENT and a message naming the limit
 *   row_ptr           rows + 1 ascending int32 offsets, row_ptr[rows] = number of nonzeros (< 2^31)
 *   col_idx, values   column (int32, in [0, cols)) and value of every nonzero, row by row
 *   v (cols), out (rows)
 *   ALL FIVE ARRAYS ARE DEVICE POINTERS on `device` (-1: the current device); host pointers fault.
 * Rows that several lanes share are summed in a lane order that may differ from pdlp_b200_matvec's
 * (the table is cut differently): results agree to rounding (1e-12 relative), not bit for bit.
 * The static array is shared by every call on a GPU and keeps the last call's contents; the call
 * clears what it needs cleared.  Calls that overlap on the same GPU (other streams, other threads)
 * race on it: serialise them. */
#define PDLP_B200_NO_ALLOC_MAX_ROWS 262144
pdlp_b200_status pdlp_b200_matvec_no_alloc(int64_t rows, int64_t cols, const int32_t* row_ptr, const int32_t* col_idx, const double* values,
                                           const double* v, double* out, int device, void* stream);

/* Host only (no device needed): internal::ComputeStatuses (primal_dual_hybrid_gradient.h:164, .cc:2835-2883), the
 * Glop basis statuses of a PDLP point, which the reference computes before it asks Glop's preprocessor to recover the
 * original solution (:1295).  For the postsolve callback of pdlp_b200_solve_presolved.  Values are Glop's
 * (lp_types.h:201-241): 0 BASIC, 1 FIXED_VALUE, 2 AT_LOWER_BOUND, 3 AT_UPPER_BOUND (4 FREE is never produced).
 * Only the four bound vectors of `problem` are read. */
pdlp_b200_status pdlp_b200_compute_statuses(const pdlp_b200_problem* problem, const double* primal_solution /*n*/, const double* dual_solution /*m*/,
                                            int8_t* variable_statuses /*n*/, int8_t* constraint_statuses /*m*/);

/* Host only (no device needed): the work-block partition the row-stream kernels run on
 * (at most 2048 nonzeros and 1024 rows per block, a longer row alone in its block),
 * from `rows + 1` compressed offsets.  cut in (0, rows): no block straddles that row (the
 * shared / private split of a sharded block), *cut_block = the block starting there;
 * cut < 0: none, *cut_block = -1.  block_start has room for rows + 1 entries; the first
 * *num_blocks + 1 are written.  For the CPU-side tests of the invariants the kernels rely on. */
pdlp_b200_status pdlp_b200_work_blocks(const int64_t* offsets, int64_t rows, int64_t cut, int32_t* block_start, int64_t* num_blocks,
                                       int32_t* cut_block, int32_t* max_row_nnz);

#ifdef __cplusplus
}
#endif
#endif /* PDLP_B200_H_ */
