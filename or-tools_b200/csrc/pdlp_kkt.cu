// Termination statistics on the device: every reduction that
// ComputeConvergenceInformation / ComputeInfeasibilityInformation
// (ortools/pdlp/iteration_stats.cc:349-518) and SetActiveSetInformation
// (primal_dual_hybrid_gradient.cc:1343-1412) need for one point, produced by ONE pass
// over CSR(K) (K x fused with the row-side reductions) and ONE pass over CSC(K)
// (K^T y fused with the column-side reductions).  The reference spends 4 SpMVs and
// ~20 vector passes on the same numbers.  Also ReducedCosts (iteration_stats.cc:533-565).
#include <algorithm>
#include <stdexcept>

#include "pdlp_launch.cuh"
#include "pdlp_kkt_kernels.cuh"

namespace pdlp_b200 {

namespace {

inline double NonNeg(double v) { return v > 0.0 ? v : 0.0; }  // max-slots start at 0 in the reference

}  // namespace

namespace {

KktTotals TotalsFromHost(const DeviceProblem& p, const double* h) {
  KktTotals t{};
  if (p.m > 0) {
    t.primal_res_sumsq = h[0]; t.primal_res_linf = NonNeg(h[1]); t.primal_res_cw = NonNeg(h[2]); t.dual_bounds_term = h[3];
    t.y_linf = NonNeg(h[4]); t.y_sumsq = h[5]; t.primal_res_linf_homogeneous = NonNeg(h[6]);
    t.active_dual_count = h[7]; t.active_dual_change = h[8];
  }
  if (p.n > 0) {
    const double* c = h + kMaxSlots;
    t.c_dot_x = c[0]; t.x_q_x = c[1]; t.correction = c[2]; t.full_correction = c[3];
    t.dual_res_linf = NonNeg(c[4]); t.dual_res_sumsq = c[5]; t.dual_res_cw = NonNeg(c[6]);
    t.x_linf = NonNeg(c[7]); t.x_sumsq = c[8]; t.ray_correction = c[9]; t.ray_dual_res_linf = NonNeg(c[10]);
    t.sign_violation = NonNeg(c[11]); t.qx_linf = NonNeg(c[12]); t.active_primal_count = c[13]; t.active_primal_change = c[14];
  }
  return t;
}

KktRowOp RowOpOf(const Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* y_start, double cw_primal_offset) {
  const SparseRows& csr = p.K.by_row;
  return KktRowOp{{}, csr.val.get(), csr.idx.get(), x, y, y_start, p.row_scale.get(), p.clb.get(), p.cub.get(), cw_primal_offset, static_cast<int>(ws.dual_skip_below)};
}
KktColOp ColOpOf(const DeviceProblem& p, const double* x, const double* y, const double* x_start, double cw_dual_offset, bool flag) {
  const SparseRows& csc = p.K.by_col;
  return KktColOp{{}, csc.val.get(), csc.idx.get(), y, x, x_start, p.col_scale.get(), p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), cw_dual_offset, flag ? 1 : 0};
}

template <int P>
void LaunchMultiDot(Workspace& ws, const SparseRows& s, const double* const* in, double* const* out) {
  MultiDotOp<P> op{};
  op.val = s.val.get();
  op.idx = s.idx.get();
  for (int k = 0; k < P; ++k) { op.in[k] = in[k]; op.out[k] = out[k]; }
  static const int resident = [] {
    int device = 0, sms = 148, per_sm = PDLP_ROWSTREAM_MIN_CTAS;
    if (cudaGetDevice(&device) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rowstream_multi_kernel<MultiDotOp<P>>, kBlock, 0) != cudaSuccess || per_sm < 1) per_sm = PDLP_ROWSTREAM_MIN_CTAS;
    return std::min(kMaxGrid, sms * per_sm);
  }();
  const int grid = std::max(1, std::min(s.num_blocks, resident));
  rowstream_multi_kernel<MultiDotOp<P>><<<grid, kBlock, 0, ws.stream>>>(op, s.View());
  ++ws.launches;
}
void MultiDot(Workspace& ws, const SparseRows& s, int count, const double* const* in, double* const* out) {
  if (count == 1) LaunchMultiDot<1>(ws, s, in, out);
  else if (count == 2) LaunchMultiDot<2>(ws, s, in, out);
  else if (count == 3) LaunchMultiDot<3>(ws, s, in, out);
}

}  // namespace

void ComputeProducts(Workspace& ws, const DeviceProblem& p, const PointProducts* points, int count) {
  if (IsSharded(ws)) throw std::logic_error("pdlp_b200: ComputeProducts is a single-GPU pass");
  if (count > 3) throw std::logic_error("pdlp_b200: at most three points per products pass");
  const double* in[3];
  double* out[3];
  int rows = 0, cols = 0;
  if (p.m > 0) {
    for (int k = 0; k < count; ++k)
      if (points[k].x != nullptr && points[k].kx != nullptr) { in[rows] = points[k].x; out[rows] = points[k].kx; ++rows; }
    MultiDot(ws, p.K.by_row, rows, in, out);
  }
  if (p.n > 0) {
    for (int k = 0; k < count; ++k)
      if (points[k].y != nullptr && points[k].kty != nullptr) { in[cols] = points[k].y; out[cols] = points[k].kty; ++cols; }
    MultiDot(ws, p.K.by_col, cols, in, out);
  }
}

KktTotals ComputeKktTotalsFromProducts(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* x_start, const double* y_start,
                                       const double* kx, const double* kty, double cw_primal_offset, double cw_dual_offset, bool flag) {
  if (p.m == 0 && p.n == 0) return KktTotals{};
  if (p.m > 0) {
    const int g = EnqueueMap(ws, KktRowMapOp{{}, RowOpOf(ws, p, x, y, y_start, cw_primal_offset), kx}, p.m, 0);
    EnqueueFinalize<KktRowMapOp>(ws, g, 0, 0, Side::kDual);
  }
  if (p.n > 0) {
    const int g = EnqueueMap(ws, KktColMapOp{{}, ColOpOf(p, x, y, x_start, cw_dual_offset, flag), kty}, p.n, 1);
    EnqueueFinalize<KktColMapOp>(ws, g, 1, kMaxSlots, Side::kSharded);
  }
  return TotalsFromHost(p, FetchTotals(ws, 2 * kMaxSlots));
}

KktTotals ComputeKktTotals(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* x_start, const double* y_start,
                           double cw_primal_offset, double cw_dual_offset, bool flag) {
  const SparseRows& csr = p.K.by_row;
  const SparseRows& csc = p.K.by_col;
  int g_row = 0, g_col = 0;
  if (p.m > 0) {
    g_row = EnqueueRows(ws, RowOpOf(ws, p, x, y, y_start, cw_primal_offset), csr, 0);
    EnqueueFinalize<KktRowOp>(ws, g_row, 0, 0, SideOfRows(csr));
  }
  if (p.n > 0) {
    g_col = EnqueueRows(ws, ColOpOf(p, x, y, x_start, cw_dual_offset, flag), csc, 1);
    EnqueueFinalize<KktColOp>(ws, g_col, 1, kMaxSlots, SideOfRows(csc));
  }
  if (p.m == 0 && p.n == 0) return KktTotals{};
  return TotalsFromHost(p, FetchTotals(ws, 2 * kMaxSlots));
}

void ComputeReducedCosts(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, bool zero_objective, bool flag, double* rc) {
  const SparseRows& csc = p.K.by_col;
  if (p.n > 0) {
    EnqueueRows(ws, ReducedCostOp{{}, {}, csc.val.get(), csc.idx.get(), y, x, p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), rc, zero_objective ? 1 : 0, flag ? 1 : 0}, csc);
  }
}

}  // namespace pdlp_b200
