// Setup-time and utility device ops: problem statistics, Ruiz / L2 rescaling, bound
// projections, plain SpMV.  Device replacements for
//   ortools/pdlp/sharded_optimization_utils.cc:83-444,701-750,
//   ortools/pdlp/sharded_quadratic_program.cc:97-175, ortools/pdlp/sharder.cc:163-335.
#include <algorithm>

#include "pdlp_launch.cuh"

namespace pdlp_b200 {

void Workspace::Init(int64_t n_, int64_t m_) {
  n = n_; m = m_;
  PDLP_CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  // Two launches may share a region's table (sharded CSR pass: private rows, then shared rows).
  for (auto& p : partials) { p.Resize(static_cast<size_t>(2 * kMaxGrid) * kMaxSlots); p.Zero(stream); }
  totals.Resize(kMaxSlots * 4);
  totals.Zero(stream);
  PDLP_CUDA_CHECK(cudaMallocHost(&host_totals, kMaxSlots * 4 * sizeof(double)));
  for (auto& v : scratch_n) v.Resize(n > 0 ? n : 1);
  for (auto& v : scratch_m) v.Resize(m > 0 ? m : 1);
  for (auto& v : scratch_nm) v.Resize(n + m > 0 ? n + m : 1);
  combine_tmp.Resize(kMaxSlots);
}

void Workspace::Destroy() {
  if (host_totals != nullptr) cudaFreeHost(host_totals);
  host_totals = nullptr;
  if (stream != nullptr) cudaStreamDestroy(stream);
  stream = nullptr;
}

void Iterate::Init(int64_t n, int64_t m) {
  for (auto& v : x) v.Resize(n > 0 ? n : 1);
  for (auto& v : kty) v.Resize(n > 0 ? n : 1);
  for (auto& v : y) v.Resize(m > 0 ? m : 1);
  xbar.Resize(n > 0 ? n : 1);
  avg_x.Resize(n > 0 ? n : 1);
  avg_y.Resize(m > 0 ? m : 1);
  x_start.Resize(n > 0 ? n : 1);
  y_start.Resize(m > 0 ? m : 1);
  state.Resize(1);
}

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

// Matrix scaling runs per row so the row factor is read once per row.
__global__ void __launch_bounds__(kBlock) scale_matrix_kernel(const int* ptr, const int* idx, double* val, int rows, const double* idx_scale,
                                                             const double* row_scale) {
  // One warp per row, grid-stride.
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * kBlock + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * kBlock) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const double rs = row_scale[r];
    for (int k = ptr[r] + lane; k < ptr[r + 1]; k += 32) val[k] *= idx_scale[idx[k]] * rs;
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

VectorInfo ComputeVectorInfo(Workspace& ws, InfoKind kind, const double* a, const double* b, int64_t count, Side side) {
  VectorInfo r;
  double t[7] = {0, 0, 0, -INFINITY, -INFINITY, 0, 0};
  if (count > 0 || IsSharded(ws)) {
    // Terms this rank owns: [skip, count) (pdlp_ops.cuh Side).
    const int64_t skip = std::min(ws.SkipBelow(side), count);
    const double* h = RunMap(ws, InfoOp{{}, a + skip, b != nullptr ? b + skip : nullptr, static_cast<int>(kind)}, count - skip, side);
    for (int i = 0; i < 7; ++i) t[i] = h[i];
  }
  r.num_infinite = static_cast<int64_t>(t[0]);
  r.num_zero = static_cast<int64_t>(t[1]);
  r.num_finite_nonzero = static_cast<int64_t>(t[2]);
  r.largest = r.num_finite_nonzero > 0 ? t[3] : 0.0;
  r.smallest = r.num_finite_nonzero > 0 ? -t[4] : 0.0;
  const int64_t denom = r.num_finite_nonzero + r.num_zero;
  r.average = denom > 0 ? t[5] / static_cast<double>(denom) : std::numeric_limits<double>::quiet_NaN();
  r.l2_norm = std::sqrt(t[6]);
  return r;
}

void ReplaceLargeBoundsWithInfinity(Workspace& ws, double* clb, double* cub, int64_t m, double threshold) {
  if (m > 0) EnqueueMap(ws, ClampInfOp{{}, {}, clb, cub, threshold}, m);
}

void ScaledRowNorms(Workspace& ws, const SparseRows& s, const double* idx_scale, const double* row_scale, bool l2, double* out) {
  if (s.rows > 0) EnqueueRows(ws, RowNormOp{{}, {}, s.val.get(), s.idx.get(), idx_scale, row_scale, out, l2 ? 1 : 0}, s);
}

void DivideBySqrtOfDivisor(Workspace& ws, const double* divisor, double* v, int64_t count) {
  if (count > 0) EnqueueMap(ws, DivSqrtOp{{}, {}, divisor, v}, count);
}

void FillVector(Workspace& ws, double* v, int64_t count, double value) {
  if (count > 0) EnqueueMap(ws, FillOp{{}, {}, v, value}, count);
}

void ScaleMatrixValues(Workspace& ws, SparseRows& s, const double* idx_scale, const double* row_scale) {
  if (s.rows == 0 || s.nnz == 0) return;
  const int64_t warps_needed = s.rows;
  int64_t grid = (warps_needed * 32 + kBlock - 1) / kBlock;
  if (grid > kMaxGrid) grid = kMaxGrid;
  scale_matrix_kernel<<<static_cast<int>(grid), kBlock, 0, ws.stream>>>(s.ptr.get(), s.idx.get(), s.val.get(), s.rows, idx_scale, row_scale);
  ++ws.launches;
}

void ScaleProblemVectors(Workspace& ws, DeviceProblem& p, const double* col_scale, const double* row_scale) {
  if (p.n > 0) EnqueueMap(ws, ScalePrimalOp{{}, {}, p.c.get(), p.vlb.get(), p.vub.get(), p.is_lp ? nullptr : p.q.get(), col_scale}, p.n);
  if (p.m > 0) EnqueueMap(ws, ScaleDualOp{{}, {}, p.clb.get(), p.cub.get(), row_scale}, p.m);
}

void ProjectToPrimalBounds(Workspace& ws, const DeviceProblem& p, double* x) {
  if (p.n > 0) EnqueueMap(ws, ProjectPrimalOp{{}, {}, x, p.vlb.get(), p.vub.get()}, p.n);
}
void ProjectToDualBounds(Workspace& ws, const DeviceProblem& p, double* y) {
  if (p.m > 0) EnqueueMap(ws, ProjectDualOp{{}, {}, y, p.clb.get(), p.cub.get()}, p.m);
}
void DivideInPlace(Workspace& ws, double* v, const double* divisor, int64_t count) {
  if (count > 0) EnqueueMap(ws, DivOp{{}, {}, v, divisor}, count);
}
void MultiplyInPlace(Workspace& ws, double* v, const double* factor, int64_t count) {
  if (count > 0) EnqueueMap(ws, MulOp{{}, {}, v, factor}, count);
}
void CopyVector(Workspace& ws, double* dst, const double* src, int64_t count) {
  if (count > 0) PDLP_CUDA_CHECK(cudaMemcpyAsync(dst, src, count * sizeof(double), cudaMemcpyDeviceToDevice, ws.stream));
}
void SubtractVectors(Workspace& ws, double* dst, const double* a, const double* b, int64_t count) {
  if (count > 0) EnqueueMap(ws, SubOp{{}, {}, dst, a, b}, count);
}
void NormAndMax(Workspace& ws, const double* v, int64_t count, double* l2, double* linf, Side side) {
  *l2 = 0.0; *linf = 0.0;
  if (count == 0 && !IsSharded(ws)) return;
  const int64_t skip = std::min(ws.SkipBelow(side), count);
  const double* h = RunMap(ws, NormMaxOp{{}, v + skip}, count - skip, side);
  *l2 = std::sqrt(h[0]);
  *linf = h[1] < 0 ? 0.0 : h[1];
}
double SquaredDistance(Workspace& ws, const double* a, const double* b, int64_t count, Side side) {
  if (count == 0 && !IsSharded(ws)) return 0.0;
  const int64_t skip = std::min(ws.SkipBelow(side), count);
  return RunMap(ws, SqDistOp{{}, a + skip, b + skip}, count - skip, side)[0];
}

void RowDot(Workspace& ws, const SparseRows& s, const double* v, double* out) {
  if (s.rows > 0) EnqueueRows(ws, RowDotOp{{}, {}, s.val.get(), s.idx.get(), v, out}, s);
}

}  // namespace pdlp_b200
