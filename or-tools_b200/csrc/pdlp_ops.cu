// Setup-time and utility device ops: problem statistics, Ruiz / L2 rescaling, bound
// projections, plain SpMV.  Device replacements for
//   ortools/pdlp/sharded_optimization_utils.cc:83-444,701-750,
//   ortools/pdlp/sharded_quadratic_program.cc:97-175, ortools/pdlp/sharder.cc:163-335.
#include <algorithm>

#include "pdlp_launch.cuh"
#include "pdlp_ops_kernels.cuh"

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
  const int grid = std::max(1, std::min(s.num_blocks, kMaxGrid));
  scale_matrix_kernel<<<grid, kBlock, 0, ws.stream>>>(s.View(), s.idx.get(), s.val.get(), idx_scale, row_scale);
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
