// Restart-decision numerics on the device: Lagrangian gradients
// (ortools/pdlp/sharded_optimization_utils.cc:446-527) and the Euclidean-norm localized
// Lagrangian bounds (ortools/pdlp/trust_region.cc:886-1016; the max-norm variant :797-884 as two
// masked sub-problems of the same code) with the joint primal-dual
// bound-constrained trust-region solve (trust_region.cc:332-449, trust_region.h:193-337).
//
// The reference finds the trust-region step by repeated median pivoting over shrinking
// index lists.  That is sequential in nature; here the same step (it is the unique root
// of a monotone function) is found by a monotone Newton iteration: with s = delta^2,
//   g(s) = sum_i a_i min(s, s_i),  a_i = obj_i^2 / w_i,  s_i = (critical step)_i^2
// is concave, increasing and piecewise linear, so Newton from the left never overshoots
// and lands exactly on the root once no further breakpoint is crossed.  Each Newton step
// is one deterministic reduction pass over the n + m critical step sizes.
#include <cfloat>

#include "pdlp_launch.cuh"
#include "pdlp_restart_kernels.cuh"

namespace pdlp_b200 {

namespace {
struct DeviceRunner {
  Workspace& ws;
  const DeviceProblem& p;
  template <class Op> const double* CsrRows(const Op& op) { return RunRows(ws, op, p.K.by_row); }
  void ColumnDot(const double* y, double* out) { RowDot(ws, p.K.by_col, y, out); }
  template <class Op> const double* Primal(const Op& op, int64_t count) { return RunMap(ws, op, count, Side::kSharded); }
  template <class Op> const double* Dual(const Op& op, int64_t count) { return RunMap(ws, op, count, Side::kDual); }
  template <class Op> const double* Joint(const Op& op, int64_t count) { return RunMap(ws, op, count, Side::kJoint); }
};
}  // namespace

LocalizedBounds ComputeLocalizedLagrangianBounds(Workspace& ws, const DeviceProblem& p, const double* x, const double* y, const double* kty_or_null,
                                                 double primal_weight, double radius, bool use_diagonal_qp_solver, double diagonal_qp_tolerance,
                                                 int* tr_passes, bool max_norm, const double* x_start, const double* y_start, const double* kx_or_null) {
  const SparseRows& csr = p.K.by_row;
  const RestartProblemView view{csr.val.get(), csr.idx.get(), p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), p.clb.get(), p.cub.get(), p.n, p.m, p.is_lp,
                                ws.dual_skip_below, ws.scratch_n[0].get(), ws.scratch_m[0].get(), ws.scratch_n[1].get(), ws.scratch_nm[0].get(),
                                ws.scratch_nm[1].get()};
  DeviceRunner run{ws, p};
  return LocalizedLagrangianBoundsWith(run, view, x, y, kty_or_null, primal_weight, radius, use_diagonal_qp_solver, diagonal_qp_tolerance, tr_passes, max_norm, x_start, y_start, kx_or_null);
}

}  // namespace pdlp_b200
