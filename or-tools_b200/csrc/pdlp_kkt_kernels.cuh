// Device side of the termination statistics (overview: pdlp_kkt.cu): the row-side and column-side
// ops of the two fused passes and the reduced-cost op.  Pure device code over pdlp_skeletons.cuh,
// included by pdlp_kkt.cu only - and, with PDLP_EMULATE, by tests/kernel_emu/ to run them on CPU.
#ifndef PDLP_B200_KKT_KERNELS_CUH_
#define PDLP_B200_KKT_KERNELS_CUH_

#include "pdlp_skeletons.cuh"

namespace pdlp_b200 {

namespace {

// iteration_stats.cc:140-151.
__device__ __forceinline__ bool HandleAsReducedCost(bool flag, double g, double x, double lb, double ub) {
  if (g == 0.0) return true;
  const double active = g > 0.0 ? lb : ub;
  if (flag) return fabs(x - active) <= fabs(x);
  return isfinite(active);
}

// Row side.  PrimalResidualNorms (iteration_stats.cc:66-134) in both its plain and
// homogeneous forms, DualObjectiveBoundsTerm (:294-316), scaled norms of y.
struct KktRowOp : AlwaysActive {
  static constexpr int kNumSlots = 9;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return (s == 1 || s == 2 || s == 4 || s == 6) ? kMax : kSum; }
  const double* val; const int* idx; const double* x;
  const double* y; const double* y_start; const double* row_scale; const double* clb; const double* cub;
  double cw_offset;
  int skip_below;  // sharded: rows below it are shared with other ranks and counted on rank 0 only
  __device__ double Load(int64_t k) const { return val[k] * x[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  __device__ void RowFinish(int i, double kx, Slots<9>& r) const {
    if (i < skip_below) return;
    const double lb = clb[i], ub = cub[i], rs = row_scale[i], yi = y[i];
    {
      double scaled_res = 0.0, res_bound = 0.0;
      if (kx > ub) { scaled_res = kx - ub; res_bound = ub; }
      else if (kx < lb) { scaled_res = lb - kx; res_bound = lb; }
      const double res = scaled_res / rs;
      r.v[0] += res * res;
      r.v[1] = fmax(r.v[1], res);
      if (res > 0.0) r.v[2] = fmax(r.v[2], res / (cw_offset + fabs(res_bound / rs)));
    }
    {
      const double ubh = isfinite(ub) ? 0.0 : ub, lbh = isfinite(lb) ? 0.0 : lb;
      double scaled_res = 0.0;
      if (kx > ubh) scaled_res = kx - ubh;
      else if (kx < lbh) scaled_res = lbh - kx;
      r.v[6] = fmax(r.v[6], scaled_res / rs);
    }
    if (yi > 0.0) r.v[3] += lb * yi;
    else if (yi < 0.0) r.v[3] += ub * yi;
    const double sy = yi * rs;
    r.v[4] = fmax(r.v[4], fabs(sy));
    r.v[5] += sy * sy;
    const bool free_row = (lb == -INFINITY && ub == INFINITY);
    const bool act = (yi != 0.0) || free_row;
    const bool act0 = (y_start[i] != 0.0) || free_row;
    r.v[7] += act ? 1.0 : 0.0;
    r.v[8] += (act != act0) ? 1.0 : 0.0;
  }
};

// Column side.  Objective pieces (:239-263), PrimalGradientFromObjectiveProduct
// (:268-289), DualResidualNorms (:163-236) for the point and for the ray (-K^T y),
// primal-ray sign violations (:472-494), scaled norms of x.
struct KktColOp : AlwaysActive {
  static constexpr int kNumSlots = 15;
  __host__ __device__ static constexpr int SlotKindOf(int s) {
    return (s == 4 || s == 6 || s == 7 || s == 10 || s == 11 || s == 12) ? kMax : kSum;
  }
  const double* val; const int* idx; const double* y;
  const double* x; const double* x_start; const double* col_scale; const double* c; const double* q; const double* vlb; const double* vub;
  double cw_offset; int flag;
  __device__ double Load(int64_t k) const { return val[k] * y[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  __device__ void RowFinish(int j, double kty, Slots<15>& r) const {
    const double xj = x[j], cj = c[j], cs = col_scale[j], lb = vlb[j], ub = vub[j];
    const double qx = (q != nullptr) ? q[j] * xj : 0.0;
    r.v[0] += cj * xj;
    r.v[1] += qx * xj;
    const double grad = qx + (cj - kty);
    if (grad != 0.0) {
      const double bound = grad > 0.0 ? lb : ub;
      r.v[3] += bound * grad;
      if (HandleAsReducedCost(flag != 0, grad, xj, lb, ub)) {
        r.v[2] += bound * grad;
      } else {
        const double res = fabs(grad) / cs;
        r.v[4] = fmax(r.v[4], res);
        r.v[5] += res * res;
        if (res > 0.0) r.v[6] = fmax(r.v[6], res / (cw_offset + fabs(cj / cs)));
      }
    }
    const double sx = xj * cs;
    r.v[7] = fmax(r.v[7], fabs(sx));
    r.v[8] += sx * sx;
    const double gray = -kty;
    if (gray != 0.0) {
      const double bound = gray > 0.0 ? lb : ub;
      if (HandleAsReducedCost(flag != 0, gray, xj, lb, ub)) r.v[9] += bound * gray;
      else r.v[10] = fmax(r.v[10], fabs(gray) / cs);
    }
    if (isfinite(lb)) r.v[11] = fmax(r.v[11], -xj * cs);
    if (isfinite(ub)) r.v[11] = fmax(r.v[11], xj * cs);
    r.v[12] = fmax(r.v[12], fabs(qx));
    const bool act = xj > lb && xj < ub;
    const double x0 = x_start[j];
    const bool act0 = x0 > lb && x0 < ub;
    r.v[13] += act ? 1.0 : 0.0;
    r.v[14] += (act != act0) ? 1.0 : 0.0;
  }
};

// Products of P vectors with one stream of the matrix (rowstream_multi_kernel): out[p][row] = sum val v[p][idx].
template <int P>
struct MultiDotOp : AlwaysActive {
  static constexpr int kPoints = P;
  const double* val; const int* idx;
  const double* in[P]; double* out[P];
  __device__ double Value(int64_t k) const { return val[k]; }
  __device__ int Index(int64_t k) const { return idx[k]; }
  __device__ double Gather(int point, int index) const { return in[point][index]; }
  __device__ void Store(int point, int row, double sum) const { out[point][row] = sum; }
};

// The two epilogues as dense passes over a product that is already there (same code, same slots).  Every thread has
// a row here, where the fused kernels run their epilogue on one lane per row: the vectors are read coalesced.
struct KktRowMapOp : AlwaysActive {
  static constexpr int kNumSlots = KktRowOp::kNumSlots;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return KktRowOp::SlotKindOf(s); }
  KktRowOp row; const double* kx;
  __device__ void Apply(int64_t i, Slots<kNumSlots>& r) const { row.RowFinish(static_cast<int>(i), kx[i], r); }
};
struct KktColMapOp : AlwaysActive {
  static constexpr int kNumSlots = KktColOp::kNumSlots;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return KktColOp::SlotKindOf(s); }
  KktColOp col; const double* kty;
  __device__ void Apply(int64_t j, Slots<kNumSlots>& r) const { col.RowFinish(static_cast<int>(j), kty[j], r); }
};

struct ReducedCostOp : AlwaysActive, NoSlots {
  const double* val; const int* idx; const double* y;
  const double* x; const double* c; const double* q; const double* vlb; const double* vub; double* rc;
  int zero_objective; int flag;
  __device__ double Load(int64_t k) const { return val[k] * y[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  __device__ void RowFinish(int j, double kty, Slots<0>&) const {
    const double xj = x[j];
    double g;
    if (zero_objective) g = -kty;
    else g = ((q != nullptr) ? q[j] * xj : 0.0) + (c[j] - kty);
    if (g != 0.0 && !HandleAsReducedCost(flag != 0, g, xj, vlb[j], vub[j])) g = 0.0;
    rc[j] = g;
  }
};

}  // namespace

}  // namespace pdlp_b200
#endif  // PDLP_B200_KKT_KERNELS_CUH_
