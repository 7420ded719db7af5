// Termination statistics on the device: every reduction that
// ComputeConvergenceInformation / ComputeInfeasibilityInformation
// (ortools/pdlp/iteration_stats.cc:349-518) and SetActiveSetInformation
// (primal_dual_hybrid_gradient.cc:1343-1412) need for one point, produced by ONE pass
// over CSR(K) (K x fused with the row-side reductions) and ONE pass over CSC(K)
// (K^T y fused with the column-side reductions).  The reference spends 4 SpMVs and
// ~20 vector passes on the same numbers.  Also ReducedCosts (iteration_stats.cc:533-565).
#include "pdlp_launch.cuh"
#include "pdlp_kkt_kernels.cuh"

namespace pdlp_b200 {

namespace {

inline double NonNeg(double v) { return v > 0.0 ? v : 0.0; }  // max-slots start at 0 in the reference

}  // namespace

KktTotals ComputeKktTotals(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* x_start, const double* y_start,
                           double cw_primal_offset, double cw_dual_offset, bool flag) {
  KktTotals t{};
  const SparseRows& csr = p.K.by_row;
  const SparseRows& csc = p.K.by_col;
  int g_row = 0, g_col = 0;
  if (p.m > 0) {
    g_row = EnqueueRows(ws, KktRowOp{{}, csr.val.get(), csr.idx.get(), x, y, y_start, p.row_scale.get(), p.clb.get(), p.cub.get(), cw_primal_offset, static_cast<int>(ws.dual_skip_below)}, csr, 0);
    EnqueueFinalize<KktRowOp>(ws, g_row, 0, 0, SideOfRows(csr));
  }
  if (p.n > 0) {
    g_col = EnqueueRows(ws, KktColOp{{}, csc.val.get(), csc.idx.get(), y, x, x_start, p.col_scale.get(), p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(),
                                     cw_dual_offset, flag ? 1 : 0}, csc, 1);
    EnqueueFinalize<KktColOp>(ws, g_col, 1, kMaxSlots, SideOfRows(csc));
  }
  if (p.m == 0 && p.n == 0) return t;
  const double* h = FetchTotals(ws, 2 * kMaxSlots);
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

void ComputeReducedCosts(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, bool zero_objective, bool flag, double* rc) {
  const SparseRows& csc = p.K.by_col;
  if (p.n > 0) {
    EnqueueRows(ws, ReducedCostOp{{}, {}, csc.val.get(), csc.idx.get(), y, x, p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), rc, zero_objective ? 1 : 0, flag ? 1 : 0}, csc);
  }
}

}  // namespace pdlp_b200
