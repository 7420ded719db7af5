// Device side and driver of the restart criteria (overview and reference lines: pdlp_restart.cu): the
// Lagrangian-gradient ops, the trust-region ops and the host loop that finds the trust-region step,
// written against a Runner so that the same code runs on the device (pdlp_restart.cu) and, with
// PDLP_EMULATE, on CPU (tests/kernel_emu/).  No CUDA runtime calls here.
#ifndef PDLP_B200_RESTART_KERNELS_CUH_
#define PDLP_B200_RESTART_KERNELS_CUH_

#include <cfloat>
#include <cmath>

#include "pdlp_restart_types.h"
#include "pdlp_skeletons.cuh"

namespace pdlp_b200 {

namespace {

// The joint problem of trust_region.cc:886-974 (JointTrustRegionProblem).
struct JointView {
  const double* gx; const double* gy; const double* x; const double* y;
  const double* vlb; const double* vub; const double* clb; const double* cub; const double* q;
  int64_t n; double w_primal, w_dual;
  int64_t dual_skip_below;  // sharded solve: dual rows below it are shared and counted on rank 0 only
  int part = 0;             // 0: the joint problem; 1 / 2: only its primal / dual entries (the two max-norm problems, trust_region.cc:797-884)
  __device__ bool InPart(int64_t i) const { return part == 0 || (part == 1) == (i < n); }
  __device__ bool Counts(int64_t i) const { return InPart(i) && (i < n || i - n >= dual_skip_below); }
  __device__ double Objective(int64_t i) const { return i < n ? gx[i] : -gy[i - n]; }
  __device__ double LowerBound(int64_t i) const { return i < n ? vlb[i] : (isfinite(cub[i - n]) ? -INFINITY : 0.0); }
  __device__ double UpperBound(int64_t i) const { return i < n ? vub[i] : (isfinite(clb[i - n]) ? INFINITY : 0.0); }
  __device__ double Center(int64_t i) const { return i < n ? x[i] : y[i - n]; }
  __device__ double Weight(int64_t i) const { return i < n ? w_primal : w_dual; }
  __device__ double Diagonal(int64_t i) const { return (q != nullptr && i < n) ? q[i] : 0.0; }
  // trust_region.h:193-228.
  __device__ double ProjectedValue(int64_t i, double step) const {
    const double full = Center(i) - step * Objective(i) / Weight(i);
    return fmin(fmax(full, LowerBound(i)), UpperBound(i));
  }
};

// sou.cc:476-500.
__device__ __forceinline__ double DualSubgradientCoefficient(double lb, double ub, double dual, double primal_product) {
  if (dual < 0.0) return ub;
  if (dual > 0.0) return lb;
  if (isfinite(lb) && isfinite(ub)) {
    if (primal_product < lb) return lb;
    if (primal_product > ub) return ub;
    return primal_product;
  }
  if (isfinite(lb)) return lb;
  if (isfinite(ub)) return ub;
  return 0.0;
}

struct OneSumSlot {
  static constexpr int kNumSlots = 1;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
};

// What the two gradient passes reduce besides the Lagrangian value.  With `crit` set they also do the work of
// CriticalOp for their half of the joint problem (the gradient entry is in a register there: no second pass over
// gradient, centre and bounds) and of the first Newton pass of the trust-region step (all sums at step size 0), and
// with a start point the squared distance to it (DistanceTraveledFromLastStart, pdhg.cc:1795-1804).
//   slot 0 value   1 squared distance to the start point   2 max |objective| (max)
//   slot 3 fixed radius^2 at step 0   4 free coefficient at step 0   5 number of entries fixed at step 0
struct GradientSlots {
  static constexpr int kNumSlots = 6;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return s == 2 ? kMax : kSum; }
};
struct GradientTotals { double value = 0, distance2 = 0, max_objective = 0, fixed_r2 = 0, free_coef = 0, num_fixed = 0; };
__host__ __device__ inline void AddCritical(double obj, double w, double center, double lb, double ub, bool counts, double* crit, double* coef, Slots<6>& r) {
  double cs = INFINITY;
  if (obj != 0.0) {
    const double dist = obj > 0.0 ? lb - center : ub - center;
    cs = -w * dist / obj;
  }
  const double a = obj * obj / w;
  *crit = cs;
  *coef = a;
  r.v[2] = fmax(r.v[2], fabs(obj));
  if (!counts) return;
  if (cs <= 0.0) { r.v[3] += a * (cs * cs); r.v[5] += 1.0; }
  else r.v[4] += a;
}

// ComputeDualGradient (sou.cc:502-527) fused with K x.
struct DualGradOp : AlwaysActive, GradientSlots {
  const double* val; const int* idx; const double* x;
  const double* y; const double* clb; const double* cub; double* gy;
  int skip_below;
  const double* y_start;       // null: no distance
  double* crit; double* coef;  // the dual half of the joint arrays (null: not computed)
  double w_dual;
  __device__ double Load(int64_t k) const { return val[k] * x[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  __device__ void RowFinish(int i, double kx, Slots<6>& r) const {
    const double lb = clb[i], ub = cub[i], yi = y[i];
    const double coeff = DualSubgradientCoefficient(lb, ub, yi, kx);
    const bool counts = i >= skip_below;
    if (counts) r.v[0] += coeff * yi;
    const double g = coeff - kx;
    gy[i] = g;
    if (y_start != nullptr && counts) { const double d = yi - y_start[i]; r.v[1] += d * d; }
    // The dual entry of the joint problem: objective -g, bounds of the dual variable (JointView).
    if (crit != nullptr) AddCritical(-g, w_dual, yi, isfinite(ub) ? -INFINITY : 0.0, isfinite(lb) ? INFINITY : 0.0, counts, crit + i, coef + i, r);
  }
};

// The same epilogue as a dense pass when K x is already there (a host event computes the products of all its points
// with one stream of the matrix, pdlp_kkt.cu:ComputeProducts).
struct DualGradMapOp : AlwaysActive, GradientSlots {
  DualGradOp row; const double* kx;
  __device__ void Apply(int64_t i, Slots<6>& r) const { row.RowFinish(static_cast<int>(i), kx[i], r); }
};

// ComputePrimalGradient (sou.cc:446-474).
__device__ __forceinline__ double PrimalGradAt(double xj, double cj, const double* q, int64_t j, double kty, double* value) {
  if (q == nullptr) {
    const double g = cj - kty;
    *value = xj * g;
    return g;
  }
  const double qx = q[j] * xj;
  const double g = cj + qx - kty;
  *value = xj * (g - 0.5 * qx);
  return g;
}
struct PrimalGradMapOp : AlwaysActive, GradientSlots {
  const double* kty; const double* x; const double* c; const double* q; double* gx;
  const double* x_start;       // null: no distance
  double* crit; double* coef;  // the primal half of the joint arrays (null: not computed)
  const double* vlb; const double* vub; double w_primal;
  __device__ void Apply(int64_t j, Slots<6>& r) const {
    double v;
    const double xj = x[j];
    const double g = PrimalGradAt(xj, c[j], q, j, kty[j], &v);
    gx[j] = g;
    r.v[0] += v;
    if (x_start != nullptr) { const double d = xj - x_start[j]; r.v[1] += d * d; }
    if (crit != nullptr) AddCritical(g, w_primal, xj, vlb[j], vub[j], true, crit + j, coef + j, r);
  }
};

// Critical step sizes and radius coefficients.  Slots: max |objective| (any nonzero?).
struct CriticalOp : AlwaysActive {
  static constexpr int kNumSlots = 1;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kMax; }
  JointView p; double* crit; double* coef;
  __device__ void Apply(int64_t i, Slots<1>& r) const {
    if (!p.InPart(i)) {  // outside the sub-problem: never fixed, no radius coefficient
      crit[i] = INFINITY;
      coef[i] = 0.0;
      return;
    }
    const double obj = p.Objective(i), w = p.Weight(i);
    double cs = INFINITY;
    if (obj != 0.0) {
      const double dist = obj > 0.0 ? p.LowerBound(i) - p.Center(i) : p.UpperBound(i) - p.Center(i);
      cs = -w * dist / obj;
    }
    crit[i] = cs;
    coef[i] = obj * obj / w;
    r.v[0] = fmax(r.v[0], fabs(obj));
  }
};

// One Newton pass evaluated at kRadiusCandidates step sizes at once (ascending): for each, the
// fixed radius^2, the free coefficient and the number of fixed entries.  Slot 3 k + {0, 1, 2}.
constexpr int kRadiusCandidates = 5;
struct RadiusPassOp : AlwaysActive {
  static constexpr int kNumSlots = 3 * kRadiusCandidates;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  const double* crit; const double* coef; double delta[kRadiusCandidates]; int64_t n; int64_t dual_skip_below;
  __device__ void Apply(int64_t i, Slots<kNumSlots>& r) const {
    if (i >= n && i - n < dual_skip_below) return;
    const double cs = crit[i], a = coef[i];
    const double fixed = a * (cs * cs);
#pragma unroll
    for (int k = 0; k < kRadiusCandidates; ++k) {
      if (cs <= delta[k]) { r.v[3 * k] += fixed; r.v[3 * k + 2] += 1.0; }
      else r.v[3 * k + 1] += a;
    }
  }
};

// primal_delta / dual_delta of trust_region.cc:951-973 for the linear solver.
struct BoundsDeltaOp : AlwaysActive {
  static constexpr int kNumSlots = 2;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  JointView p; double step;
  __device__ void Apply(int64_t i, Slots<2>& r) const {
    if (!p.Counts(i)) return;
    const double sol = p.ProjectedValue(i, step);
    if (i < p.n) r.v[0] += p.gx[i] * (sol - p.x[i]);
    else r.v[1] += p.gy[i - p.n] * (sol - p.y[i - p.n]);
  }
};

// The same two sums from the critical step sizes alone: an entry fixed at its bound (crit <= step) moves by
// bound - centre = -crit objective / w, a free one by -step objective / w, so objective x movement is
// -coef min(step, crit) either way; for a dual entry the objective is -gy, hence the sign.  Reads two arrays
// instead of gradient, centre and both bounds.
struct RadiusDeltaOp : AlwaysActive {
  static constexpr int kNumSlots = 2;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  const double* crit; const double* coef; double step; int64_t n; int64_t dual_skip_below;
  __device__ void Apply(int64_t i, Slots<2>& r) const {
    if (i >= n && i - n < dual_skip_below) return;
    const double a = coef[i];
    if (a == 0.0) return;  // objective entry 0: no movement (and crit = inf)
    const double t = a * fmin(step, crit[i]);
    if (i < n) r.v[0] -= t;
    else r.v[1] += t;
  }
};

// Diagonal-QP solver pieces (trust_region.cc:615-753).
__device__ __forceinline__ double ProjectedScaledDifference(const JointView& p, int64_t i, double scaling) {
  const double w = p.Weight(i);
  return fmin(fmax((-p.Objective(i) / sqrt(w)) / (p.Diagonal(i) / w + scaling), sqrt(w) * (p.LowerBound(i) - p.Center(i))),
              sqrt(w) * (p.UpperBound(i) - p.Center(i)));
}
struct DeltaNormOp : AlwaysActive, OneSumSlot {
  JointView p; double scaling;
  __device__ void Apply(int64_t i, Slots<1>& r) const {
    if (!p.Counts(i)) return;
    const double v = ProjectedScaledDifference(p, i, scaling);
    r.v[0] += v * v;
  }
};
struct DiagonalBoundsDeltaOp : AlwaysActive {
  static constexpr int kNumSlots = 3;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  JointView p; double scaling; int at_center;
  __device__ void Apply(int64_t i, Slots<3>& r) const {
    if (!p.Counts(i)) return;
    const double c = p.Center(i);
    const double sol = at_center ? c : c + sqrt(1.0 / p.Weight(i)) * ProjectedScaledDifference(p, i, scaling);
    const double d = sol - c;
    if (i < p.n) {
      r.v[0] += p.gx[i] * d;
      if (p.q != nullptr) r.v[2] += 0.5 * p.q[i] * (d * d);
    } else {
      r.v[1] += p.gy[i - p.n] * d;
    }
  }
};

}  // namespace

// What the bounds computation needs of the device problem and its scratch space.
struct RestartProblemView {
  const double* csr_val; const int* csr_idx;
  const double* c; const double* q;  // q null for an LP
  const double* vlb; const double* vub; const double* clb; const double* cub;
  int64_t n, m; bool is_lp;
  int64_t dual_skip_below;
  double* gx; double* gy; double* kty_scratch; double* crit; double* coef;  // n, m, n, n + m, n + m doubles
};

// Runner: launches an op, finalizes its reduction slots and returns the host-visible totals.
//   const double* CsrRows(op)                 row-stream pass over the CSR copy
//   void ColumnDot(y, out)                    out = K^T y (row-stream pass over the CSC copy)
//   const double* Primal(op, count)           dense pass over this rank's columns
//   const double* Dual(op, count)             dense pass over this rank's rows
//   const double* Joint(op, count)            dense pass over (columns, rows)
// x_start / y_start (both or neither): the radius is the weighted distance of (x, y) from that point
// (DistanceTraveledFromLastStart, pdhg.cc:1795-1804), reduced inside the two gradient passes; `radius` is then ignored.
template <class Runner>
LocalizedBounds LocalizedLagrangianBoundsWith(Runner& run, const RestartProblemView& p, const double* x, const double* y, const double* kty_or_null,
                                              double primal_weight, double radius, bool use_diagonal_qp_solver, double diagonal_qp_tolerance,
                                              int* tr_passes, bool max_norm = false, const double* x_start = nullptr, const double* y_start = nullptr,
                                              const double* kx_or_null = nullptr) {
  double* gx = p.gx;
  double* gy = p.gy;
  // The default solver (joint problem, linear-time root search) takes its critical step sizes from the gradient passes.
  const bool fused = !max_norm && !use_diagonal_qp_solver;
  auto totals_of = [](const double* h) { return GradientTotals{h[0], h[1], h[2], h[3], h[4], h[5]}; };
  GradientTotals primal, dual;
  if (p.m > 0) {
    const DualGradOp row{{}, {}, p.csr_val, p.csr_idx, x, y, p.clb, p.cub, gy, static_cast<int>(p.dual_skip_below), y_start,
                         fused ? p.crit + p.n : nullptr, fused ? p.coef + p.n : nullptr, 0.5 / primal_weight};
    dual = totals_of(kx_or_null != nullptr ? run.Dual(DualGradMapOp{{}, {}, row, kx_or_null}, p.m) : run.CsrRows(row));
  }
  if (p.n > 0) {
    // When K^T y is not cached it is computed first; the gradient and its value then come
    // from the same kernel either way, so "bounds at the average" and "bounds at the
    // current point" are bit-identical whenever the two points are (the reference's
    // strict-inequality tie-break in AverageHasBetterPotential relies on that).
    const double* kty = kty_or_null;
    if (kty == nullptr) {
      run.ColumnDot(y, p.kty_scratch);
      kty = p.kty_scratch;
    }
    primal = totals_of(run.Primal(PrimalGradMapOp{{}, {}, kty, x, p.c, p.q, gx, x_start, fused ? p.crit : nullptr, fused ? p.coef : nullptr, p.vlb, p.vub, 0.5 * primal_weight}, p.n));
  }
  const double primal_value = primal.value, dual_value = dual.value;
  if (x_start != nullptr || y_start != nullptr) radius = std::sqrt((0.5 * primal_weight) * primal.distance2 + (0.5 / primal_weight) * dual.distance2);
  const double lagrangian = primal_value + dual_value;
  const int64_t total = p.n + p.m;
  JointView view{gx, gy, x, y, p.vlb, p.vub, p.clb, p.cub, p.q, p.n, 0.5 * primal_weight, 0.5 / primal_weight, p.dual_skip_below};
  int passes = 0;
  LocalizedBounds out{lagrangian, lagrangian, lagrangian, radius};
  if (total == 0) { if (tr_passes) *tr_passes = 0; return out; }

  if (use_diagonal_qp_solver && !max_norm) {  // the max-norm bounds always use the linear-time solver (trust_region.cc:1003-1007)
    double primal_delta = 0.0, dual_delta = 0.0;
    if (radius == 0.0) {
      // Solution is the centre point: both deltas are exactly zero.
    } else {
      auto norm_at = [&](double scaling) { ++passes; return std::sqrt(run.Joint(DeltaNormOp{{}, {}, view, scaling}, total)[0]); };
      double lo = 0.0, hi = 1.0;  // FindScalingFactor, trust_region.cc:683-706
      while (norm_at(hi) >= radius) { lo = hi; hi *= 2; }
      while ((hi - lo) >= diagonal_qp_tolerance * std::max(1.0, lo)) {
        const double mid = (lo + hi) / 2.0;
        if (norm_at(mid) <= radius) hi = mid; else lo = mid;
      }
      const double scaling = (hi + lo) / 2.0;
      const double* h = run.Joint(DiagonalBoundsDeltaOp{{}, view, scaling, 0}, total);
      primal_delta = h[0] + (p.is_lp ? 0.0 : h[2]);
      dual_delta = h[1];
    }
    out.lower_bound = lagrangian + primal_delta;
    out.upper_bound = lagrangian + dual_delta;
    if (tr_passes) *tr_passes = passes;
    return out;
  }

  // SolveTrustRegionStepSize (trust_region.cc:332-449) for a view of the joint index range.  `first` (the fused path):
  // the critical step sizes are already in p.crit / p.coef and the sums at step size 0 are known.
  auto solve_step = [&](const JointView& v, double target_radius, const GradientTotals* first) -> double {
    double step = 0.0;
    if (!(target_radius > 0.0)) return step;
    double* crit = p.crit;
    double* coef = p.coef;
    const double max_obj = first != nullptr ? first->max_objective : run.Joint(CriticalOp{{}, v, crit, coef}, total)[0];
    if (!(max_obj > 0.0)) return step;  // every objective entry is zero: step 0 (trust_region.cc:349-358)
    // Root of phi(d) = fixed_r2(d) + d^2 free_coef(d) - r^2 (concave piecewise in d^2) by Newton from
    // the left: from a point left of the root the Newton step lands left of the root again, and
    // once a step crosses no breakpoint it is the exact root of the final segment.  Each pass also
    // evaluates phi at four points further right (the Newton point plus 1, 3, 7, 15 times the
    // last increment); the iteration continues from the rightmost one that is still left of the
    // root.  The final value depends only on the final segment's sums, so it is the same root, bit
    // for bit, as plain Newton's - in a third of the passes where the active set changes slowly.
    const double r2 = target_radius * target_radius;
    double delta = 0.0, increment = 0.0;
    double prev_fixed = -1.0;
    bool have_first = first != nullptr;
    for (;;) {
      double at, fixed_r2, free_coef, num_fixed;
      int pick = 0;
      if (have_first) {  // the pass at step size 0 came with the gradients
        have_first = false;
        at = 0.0; fixed_r2 = first->fixed_r2; free_coef = first->free_coef; num_fixed = first->num_fixed;
      } else {
        RadiusPassOp op{{}, crit, coef, {}, p.n, v.dual_skip_below};
        static const double kAhead[kRadiusCandidates] = {0.0, 1.0, 3.0, 7.0, 15.0};
        for (int k = 0; k < kRadiusCandidates; ++k) op.delta[k] = delta + kAhead[k] * increment;
        const double* h = run.Joint(op, total);
        ++passes;
        // Rightmost candidate that is not right of the root (candidate 0 never is).
        for (int k = 1; k < kRadiusCandidates && increment > 0.0; ++k) {
          const double d = op.delta[k];
          const double phi = h[3 * k] + (d * d) * h[3 * k + 1] - r2;
          if (phi <= 0.0 && d > op.delta[pick]) pick = k;
        }
        at = op.delta[pick];
        fixed_r2 = h[3 * pick]; free_coef = h[3 * pick + 1]; num_fixed = h[3 * pick + 2];
      }
      if (!(free_coef > 0.0)) { step = DBL_MAX; break; }  // trust_region.cc:438-446
      // No breakpoint crossed since the last pass: the Newton point is the exact root.
      if (pick == 0 && num_fixed == prev_fixed) { step = at; break; }
      const double next = std::sqrt(std::max(0.0, r2 - fixed_r2) / free_coef);
      if (!(next > at)) { step = at; break; }  // already at the root up to rounding
      if (passes >= 200) { step = next; break; }
      prev_fixed = num_fixed;
      increment = next - delta;
      delta = next;
    }
    return step;
  };
  double primal_delta = 0.0, dual_delta = 0.0;
  if (max_norm) {
    // ComputeMaxNormLocalizedLagrangianBounds (trust_region.cc:797-884): the primal and the dual problem are
    // solved separately, each in the unweighted Euclidean norm, with radii sqrt(2) r / sqrt(w) and sqrt(2) r sqrt(w);
    // lower = L + primal objective value, upper = L - dual objective value (its objective is -gradient).
    JointView primal_view = view, dual_view = view;
    primal_view.part = 1; primal_view.w_primal = 1.0; primal_view.w_dual = 1.0;
    dual_view.part = 2; dual_view.w_primal = 1.0; dual_view.w_dual = 1.0;
    if (p.n > 0) {
      const double step = solve_step(primal_view, std::sqrt(2.0) * radius / std::sqrt(primal_weight), nullptr);
      primal_delta = run.Joint(BoundsDeltaOp{{}, primal_view, step}, total)[0];
    }
    if (p.m > 0) {
      const double step = solve_step(dual_view, std::sqrt(2.0) * radius * std::sqrt(primal_weight), nullptr);
      dual_delta = run.Joint(BoundsDeltaOp{{}, dual_view, step}, total)[1];
    }
  } else {
    // Shared rows of a sharded solve are replicated: max over ranks took them once, the sums counted them on rank 0.
    GradientTotals first = primal;
    first.max_objective = std::max(primal.max_objective, dual.max_objective);
    first.fixed_r2 += dual.fixed_r2; first.free_coef += dual.free_coef; first.num_fixed += dual.num_fixed;
    const double step = solve_step(view, radius, &first);
    const double* h = run.Joint(RadiusDeltaOp{{}, p.crit, p.coef, step, p.n, p.dual_skip_below}, total);
    primal_delta = h[0];
    dual_delta = h[1];
  }
  out.lower_bound = lagrangian + primal_delta;
  out.upper_bound = lagrangian + dual_delta;
  if (tr_passes) *tr_passes = passes;
  return out;
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_RESTART_KERNELS_CUH_
