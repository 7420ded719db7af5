// Device side of the setup and vector helpers (overview: pdlp_ops.cu): problem statistics, bound
// clamping, scaled row / column norms and rescaling, projections, small vector ops, the plain product.
// Pure device code over pdlp_skeletons.cuh, included by pdlp_ops.cu only - and, with PDLP_EMULATE, by
// tests/kernel_emu/ to run them on CPU.
#ifndef PDLP_B200_OPS_KERNELS_CUH_
#define PDLP_B200_OPS_KERNELS_CUH_

#include "pdlp_skeletons.cuh"

namespace pdlp_b200 {

namespace {

// ---- VectorInfoAccumulator (sou.cc:117-177) ----
struct InfoOp : AlwaysActive {
  static constexpr int kNumSlots = 7;  // ninf, nzero, nfn, max, -min, sum, sumsq
  __host__ __device__ static constexpr int SlotKindOf(int s) { return (s == 3 || s == 4) ? kMax : kSum; }
  const double* a; const double* b; int kind;
  __device__ void Apply(int64_t i, Slots<7>& acc) const {
    double v;
    if (kind == 0) {
      v = a[i];
    } else if (kind == 1) {  // CombineBounds (sou.cc:83-92)
      const double v1 = a[i], v2 = b[i];
      v = 0.0;
      if (fabs(v1) < INFINITY) v = fabs(v1);
      if (fabs(v2) < INFINITY) v = fmax(v, fabs(v2));
    } else {
      v = b[i] - a[i];
    }
    if (isinf(v)) { acc.v[0] += 1.0; }
    else if (v == 0.0) { acc.v[1] += 1.0; }
    else {
      // NaN lands here, as in the reference (its l2 norm then reports the NaN).
      const double ab = fabs(v);
      acc.v[2] += 1.0;
      acc.v[3] = fmax(acc.v[3], ab);
      acc.v[4] = fmax(acc.v[4], -ab);
      acc.v[5] += ab;
      acc.v[6] += ab * ab;
    }
  }
};

struct ClampInfOp : AlwaysActive, NoSlots {
  double* clb; double* cub; double threshold;
  __device__ void Apply(int64_t i, Slots<0>&) const {
    double l = clb[i], u = cub[i];
    if (l <= -threshold) l = -INFINITY;
    if (l >= threshold) l = INFINITY;
    if (u <= -threshold) u = -INFINITY;
    if (u >= threshold) u = INFINITY;
    clb[i] = l; cub[i] = u;
  }
};

// Scaled row norms (sharder.cc:291-335): staged value val * idx_scale[idx]; the row's own
// factor multiplies the finished norm, exactly as the reference does.
struct RowNormOp : AlwaysActive, NoSlots {
  const double* val; const int* idx; const double* idx_scale; const double* row_scale; double* out; int l2;
  __device__ double Load(int64_t k) const { return val[k] * idx_scale[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return l2 ? acc + staged * staged : fmax(acc, fabs(staged)); }
  __device__ double RowMerge(double a, double b) const { return l2 ? a + b : fmax(a, b); }
  __device__ void RowFinish(int row, double acc, Slots<0>&) const { out[row] = (l2 ? sqrt(acc) : acc) * fabs(row_scale[row]); }
  __host__ __device__ bool RowCombineMax() const { return l2 == 0; }
};

struct DivSqrtOp : AlwaysActive, NoSlots {
  const double* divisor; double* v;
  __device__ void Apply(int64_t i, Slots<0>&) const {
    const double d = divisor[i];
    if (d != 0.0) v[i] = v[i] / sqrt(d);
  }
};

struct FillOp : AlwaysActive, NoSlots {
  double* v; double value;
  __device__ void Apply(int64_t i, Slots<0>&) const { v[i] = value; }
};

// Matrix scaling over the work blocks of the row-stream kernel: a CTA takes a block (<= kStreamChunk nonzeros unless
// it is one long row), keeps the block's row offsets in shared memory and gives every THREAD a nonzero, whose row it
// finds by bisection in those offsets - the value stream is read and written fully coalesced whatever the row
// lengths are.  (A warp per row, the first version, ran at 8 % of the HBM peak on the CSC copy of C3, whose
// "rows" hold three nonzeros: profiles/r01_ncu_every_kernel_c3_v10.md.)
__global__ void __launch_bounds__(kBlock) scale_matrix_kernel(RowBlocks rb, const int* idx, double* val, const double* idx_scale, const double* row_scale) {
  __shared__ int row_raw[kMaxRowsPerBlock + 1];
  for (int b = blockIdx.x; b < rb.num_blocks; b += gridDim.x) {
    const int r0 = rb.block_start[b], nrows = rb.block_start[b + 1] - r0;
    const int p0 = rb.block_nz[b], p1 = rb.block_nz[b + 1];
    for (int i = threadIdx.x; i <= nrows; i += kBlock) row_raw[i] = rb.row_ptr[r0 + i];
    __syncthreads();
    for (int k = p0 + threadIdx.x; k < p1; k += kBlock) {
      // The row of nonzero k: the last one whose first nonzero is <= k (empty rows share their successor's offset).
      int lo = 0, hi = nrows;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (row_raw[mid] <= k) lo = mid; else hi = mid;
      }
      val[k] *= idx_scale[idx[k]] * row_scale[r0 + lo];
    }
    __syncthreads();
  }
}

struct ScalePrimalOp : AlwaysActive, NoSlots {
  double* c; double* lb; double* ub; double* q; const double* col_scale;
  __device__ void Apply(int64_t j, Slots<0>&) const {
    const double s = col_scale[j];
    c[j] = c[j] * s;
    lb[j] = lb[j] / s;
    ub[j] = ub[j] / s;
    if (q != nullptr) q[j] = q[j] * (s * s);
  }
};
struct ScaleDualOp : AlwaysActive, NoSlots {
  double* lb; double* ub; const double* row_scale;
  __device__ void Apply(int64_t i, Slots<0>&) const {
    const double s = row_scale[i];
    lb[i] = lb[i] * s;
    ub[i] = ub[i] * s;
  }
};

struct ProjectPrimalOp : AlwaysActive, NoSlots {
  double* x; const double* lb; const double* ub;
  __device__ void Apply(int64_t j, Slots<0>&) const { x[j] = fmax(fmin(x[j], ub[j]), lb[j]); }
};
struct ProjectDualOp : AlwaysActive, NoSlots {
  double* y; const double* clb; const double* cub;
  __device__ void Apply(int64_t i, Slots<0>&) const {
    double v = y[i];
    if (!isfinite(cub[i])) v = fmax(v, 0.0);
    if (!isfinite(clb[i])) v = fmin(v, 0.0);
    y[i] = v;
  }
};
struct DivOp : AlwaysActive, NoSlots {
  double* v; const double* d;
  __device__ void Apply(int64_t i, Slots<0>&) const { v[i] = v[i] / d[i]; }
};
struct MulOp : AlwaysActive, NoSlots {
  double* v; const double* f;
  __device__ void Apply(int64_t i, Slots<0>&) const { v[i] = v[i] * f[i]; }
};
struct SubOp : AlwaysActive, NoSlots {
  double* dst; const double* a; const double* b;
  __device__ void Apply(int64_t i, Slots<0>&) const { dst[i] = a[i] - b[i]; }
};
struct NormMaxOp : AlwaysActive {
  static constexpr int kNumSlots = 2;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return s == 1 ? kMax : kSum; }
  const double* v;
  __device__ void Apply(int64_t i, Slots<2>& acc) const {
    const double t = v[i];
    acc.v[0] += t * t;
    acc.v[1] = fmax(acc.v[1], fabs(t));
  }
};
struct SqDistOp : AlwaysActive {
  static constexpr int kNumSlots = 1;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  const double* a; const double* b;
  __device__ void Apply(int64_t i, Slots<1>& acc) const {
    const double d = a[i] - b[i];
    acc.v[0] += d * d;
  }
};

struct RowDotOp : AlwaysActive, NoSlots {
  const double* val; const int* idx; const double* v; double* out;
  __device__ double Load(int64_t k) const { return val[k] * v[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  __device__ void RowFinish(int row, double acc, Slots<0>&) const { out[row] = acc; }
};

}  // namespace

}  // namespace pdlp_b200
#endif  // PDLP_B200_OPS_KERNELS_CUH_
