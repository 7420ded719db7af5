// Host-callable launchers for every device operation of the PDHG path, plus the
// device-resident problem / iterate / step-state records they work on.  Each launcher
// cites the reference function it replaces (files under /root/reference/ortools/pdlp/).
#ifndef PDLP_B200_OPS_CUH_
#define PDLP_B200_OPS_CUH_

#include <cmath>
#include <limits>

#include "pdlp_comm.h"
#include "pdlp_matrix.cuh"
#include "pdlp_restart_types.h"
#include "pdlp_step_state.h"

#include <nvtx3/nvToolsExt.h>  // header-only: ranges cost nothing unless a profiler has injected its library

namespace pdlp_b200 {

// The (rescaled) problem in HBM.  Primal-side vectors have length n, dual-side m.
struct DeviceProblem {
  int64_t n = 0, m = 0;
  DeviceMatrix K;
  DVec c, q;             // q empty for an LP (diagonal of the objective matrix otherwise)
  DVec vlb, vub;         // variable bounds
  DVec clb, cub;         // constraint bounds
  DVec col_scale, row_scale;
  bool is_lp = true;
  double objective_offset = 0.0, objective_scaling_factor = 1.0;
  const double* q_or_null() const { return is_lp ? nullptr : q.get(); }
};

// Which side of a column-sharded solve a reduction's index range lives on (pdlp_launch.cuh):
//   kSharded  primal-side (this rank's columns): every entry counts, totals are all-reduced
//   kDual     dual-side (this rank's rows): the leading `shared_rows` rows exist on every rank and
//             count on rank 0 only, the rest are private to the rank; totals are all-reduced
//   kJoint    the concatenation (primal block, dual rows) of the trust-region ops
enum class Side { kDual, kSharded, kJoint };

// NVTX range over a phase of the solve (tracing): set-up stages, trial-step batches, the three parts of a host
// event.  `nsys` / `ncu --nvtx` group the kernels by them.
struct TraceRange {
  explicit TraceRange(const char* name) { nvtxRangePushA(name); }
  ~TraceRange() { nvtxRangePop(); }
  TraceRange(const TraceRange&) = delete;
  TraceRange& operator=(const TraceRange&) = delete;
};

// Launch context: stream, per-kernel partial buffers, totals, pinned read-back.
struct Workspace {
  cudaStream_t stream = nullptr;
  DVec partials[3];          // 2 * kMaxGrid * kMaxSlots each
  DVec totals;               // kMaxSlots * 4
  double* host_totals = nullptr;  // pinned, kMaxSlots * 4
  DVec scratch_n[3], scratch_m[3], scratch_nm[2];  // sized n, m, n+m on demand
  int64_t n = 0, m = 0;
  long long launches = 0;    // kernels launched (bench "gpu_launches")
  std::vector<cudaEvent_t>* profile = nullptr;  // when set, the trial-step launcher records an event around every kernel
  Comm* comm = nullptr;      // null or world 1: single GPU.  Otherwise this rank owns a column block of K
  double* exchange_rows = nullptr;  // sharded: the buffer the row sums are written to and all-reduced in (Comm::AllocateExchangeBuffer)
  double* exchange_sums = nullptr;  // sharded: 16 doubles for the step-test sums
  bool peer = false;         // sharded: row sums and step-test sums travel through peer memory (pdlp_peer.cuh) instead of ncclAllReduce
  DVec combine_tmp;          // kMaxSlots: max-reduced copy of mixed sum/max totals
  // Sharded solves: rows [0, shared_rows) of this rank's dual side are shared with other ranks
  // (their row sums are all-reduced, their values replicated), the rest touch this rank's columns
  // only.  dual_skip_below = first dual row whose reduction terms count on this rank (0 on rank 0
  // and on a single GPU, shared_rows elsewhere).
  int64_t shared_rows = 0, dual_skip_below = 0;
  // One GPU, a solve with an interrupt flag: the host-mapped copy of the flag that the decision kernel reads (pdlp_step_state.h).
  // The enqueue loops stop queueing trial steps once it is up.  Null otherwise.
  const volatile int* stop_enqueue = nullptr;
  // PDLP_B200_DEBUG_TIMING under pdlp_b200_trial_steps_profiled: events around the three launches of a peer-mode CSR pass.
  std::vector<cudaEvent_t>* profile_peer = nullptr;
  int64_t SkipBelow(Side side) const { return side == Side::kDual ? dual_skip_below : 0; }
  void Init(int64_t n_, int64_t m_);
  void Destroy();
  void Sync() { PDLP_CUDA_CHECK(cudaStreamSynchronize(stream)); }
};

// Iterate storage: double-buffered x, y, K^T y so an accepted trial is a flip of `cur`.
struct Iterate {
  DVec x[2], y[2], kty[2];
  DVec xbar;
  DVec avg_x, avg_y;
  DVec x_start, y_start;       // last restart point
  DeviceArray<StepState> state;
  void Init(int64_t n, int64_t m);
};

// ---- setup (one-time per solve) ----
struct VectorInfo {  // sharded_optimization_utils.cc:117-177 VectorInfoAccumulator
  int64_t num_finite_nonzero = 0, num_infinite = 0, num_zero = 0;
  double largest = 0, smallest = 0, average = 0, l2_norm = 0;
};
enum class InfoKind { kPlain, kCombinedBounds, kGap };
// kPlain: a[i]; kCombinedBounds: CombineBounds(a[i], b[i]) (sou.cc:83-92); kGap: b[i]-a[i].
VectorInfo ComputeVectorInfo(Workspace& ws, InfoKind kind, const double* a, const double* b, int64_t count, Side side);

// sharded_quadratic_program.cc:119-130,169-175.
void ReplaceLargeBoundsWithInfinity(Workspace& ws, double* clb, double* cub, int64_t m, double threshold);

// Norms of the compressed rows of diag(idx_scale) M diag(row_scale) (sharder.cc:291-335):
// l2 == false -> Linf, true -> L2.  out has s.rows entries.
void ScaledRowNorms(Workspace& ws, const SparseRows& s, const double* idx_scale, const double* row_scale, bool l2, double* out);
// v[i] /= sqrt(divisor[i]) where divisor[i] != 0 (sou.cc:354-365).
void DivideBySqrtOfDivisor(Workspace& ws, const double* divisor, double* v, int64_t count);
void FillVector(Workspace& ws, double* v, int64_t count, double value);
// val[k] *= idx_scale[idx[k]] * row_scale[row] (sharded_quadratic_program.cc:97-117).
void ScaleMatrixValues(Workspace& ws, SparseRows& s, const double* idx_scale, const double* row_scale);
// sharded_quadratic_program.cc:134-167 vector part.
void ScaleProblemVectors(Workspace& ws, DeviceProblem& p, const double* col_scale, const double* row_scale);
// sou.cc:721-750.
void ProjectToPrimalBounds(Workspace& ws, const DeviceProblem& p, double* x);
void ProjectToDualBounds(Workspace& ws, const DeviceProblem& p, double* y);
void DivideInPlace(Workspace& ws, double* v, const double* divisor, int64_t count);
void MultiplyInPlace(Workspace& ws, double* v, const double* factor, int64_t count);
void CopyVector(Workspace& ws, double* dst, const double* src, int64_t count);
// dst = a - b.
void SubtractVectors(Workspace& ws, double* dst, const double* a, const double* b, int64_t count);
// NaN / magnitude screen of an initial solution: out = {sum of squares, max |v|}.
void NormAndMax(Workspace& ws, const double* v, int64_t count, double* l2, double* linf, Side side);

// out = M v by rows (K^T y with the CSC copy, K x with the CSR copy); sharder.cc:163-176.
void RowDot(Workspace& ws, const SparseRows& s, const double* v, double* out);

// ---- the iteration (primal_dual_hybrid_gradient.cc:2295-2376 and :2378-2406) ----
// Enqueues `trials` trial steps of the adaptive (or constant) rule; no host sync.
void EnqueueAdaptiveTrials(Workspace& ws, const DeviceProblem& p, Iterate& it, int trials, bool constant_step);
// Malitsky-Pock (primal_dual_hybrid_gradient.cc:2211-2293): `trials` dual trials.
void EnqueueMalitskyPockTrials(Workspace& ws, const DeviceProblem& p, Iterate& it, int trials);

// ---- termination statistics (iteration_stats.cc:66-518, pdhg.cc:1343-1412) ----
struct KktTotals {
  // dual side (rows)
  double primal_res_sumsq, primal_res_linf, primal_res_cw, dual_bounds_term, y_linf, y_sumsq, primal_res_linf_homogeneous;
  double active_dual_count, active_dual_change;
  // primal side (columns)
  double c_dot_x, x_q_x, correction, full_correction, dual_res_linf, dual_res_sumsq, dual_res_cw, x_linf, x_sumsq;
  double ray_correction, ray_dual_res_linf, sign_violation, qx_linf, active_primal_count, active_primal_change;
};
// One K x pass and one K^T y pass produce every reduction ComputeConvergenceInformation
// and ComputeInfeasibilityInformation need for the point (x, y).
KktTotals ComputeKktTotals(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* x_start, const double* y_start,
                           double cw_primal_offset, double cw_dual_offset, bool handle_some_primal_gradients_on_finite_bounds_as_residuals);
// K x and K^T y of up to three points with ONE stream of each matrix copy (rowstream_multi_kernel): what a host
// event needs of the current iterate, the average and the last difference at once.  Every sum is bit-identical to
// the one a single-point pass forms.  One GPU only (a sharded CSR pass needs its exchange).
struct PointProducts {
  const double* x = nullptr;  // n entries (null: no K x wanted for this point)
  const double* y = nullptr;  // m entries (null: no K^T y wanted)
  double* kx = nullptr;       // out, m entries
  double* kty = nullptr;      // out, n entries
};
void ComputeProducts(Workspace& ws, const DeviceProblem& p, const PointProducts* points, int count);
// ComputeKktTotals from products that are already there: two dense passes.
KktTotals ComputeKktTotalsFromProducts(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* x_start, const double* y_start,
                                       const double* kx, const double* kty, double cw_primal_offset, double cw_dual_offset,
                                       bool handle_some_primal_gradients_on_finite_bounds_as_residuals);
// iteration_stats.cc:533-565; rc must not alias x.
void ComputeReducedCosts(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, bool zero_objective,
                         bool handle_some_primal_gradients_on_finite_bounds_as_residuals, double* rc);

// ---- restarts (trust_region.cc:886-1016, sou.cc:446-527) ----
double SquaredDistance(Workspace& ws, const double* a, const double* b, int64_t count, Side side);
// Euclidean-norm localized Lagrangian bounds at (x, y).  kty may be null (computed).  With a start point the radius is the
// weighted distance of (x, y) from it (pdhg.cc:1795-1804), reduced inside the gradient passes, and `radius` is ignored.
LocalizedBounds ComputeLocalizedLagrangianBounds(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* kty_or_null,
                                                 double primal_weight, double radius, bool use_diagonal_qp_solver, double diagonal_qp_tolerance,
                                                 int* tr_passes, bool max_norm = false, const double* x_start = nullptr, const double* y_start = nullptr,
                                                 const double* kx_or_null = nullptr);  // K x when the caller has it: no pass over the matrix

// K v from caller-owned device CSR arrays with the work-block table in a static device array
// (pdlp_noalloc.cu; the kernel side of pdlp_b200_matvec_no_alloc, which has no counterpart in the reference).
void MatVecNoAlloc(int64_t rows, int64_t cols, const int* row_ptr, const int* col_idx, const double* values, const double* v, double* out, cudaStream_t stream);

}  // namespace pdlp_b200
#endif  // PDLP_B200_OPS_CUH_
