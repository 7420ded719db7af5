// Device plumbing shared by every PDLP kernel file: error checking, RAII device buffers,
// deterministic block reductions, and the two kernel skeletons every op is an instance of:
//
//   map_reduce_kernel<Op>   one grid-stride pass over a dense index range [0, count)
//   rowstream_kernel<Op>    one pass over a compressed sparse matrix by rows ("row" = a
//                           column of the CSC copy or a row of the CSR copy), the device
//                           replacement of TransposedMatrixVectorProduct
//                           (ortools/pdlp/sharder.cc:163-176) and of the scaled column
//                           norms (sharder.cc:291-335).
//
// Both skeletons end with a fixed-shape two-stage reduction (thread -> warp shuffle ->
// shared memory -> one partial per block, then a finalize pass in fixed order), so
// results are bit-reproducible for a given launch configuration: the device analogue of
// the reference's "same num_shards => same roundoff" promise (solvers.proto:273-282).
#ifndef PDLP_B200_COMMON_CUH_
#define PDLP_B200_COMMON_CUH_

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <utility>

namespace pdlp_b200 {

#define PDLP_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t err__ = (expr);                                                            \
    if (err__ != cudaSuccess) {                                                            \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err__) +  \
                               " at " + __FILE__ + ":" + std::to_string(__LINE__));       \
    }                                                                                      \
  } while (0)

// Owning device array.  Allocation is synchronous (cudaMalloc); all solver buffers are
// created once per solve, never inside the iteration loop.
template <typename T>
class DeviceArray {
 public:
  DeviceArray() = default;
  explicit DeviceArray(size_t n) { Resize(n); }
  DeviceArray(const DeviceArray&) = delete;
  DeviceArray& operator=(const DeviceArray&) = delete;
  DeviceArray(DeviceArray&& o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr; o.n_ = 0; }
  DeviceArray& operator=(DeviceArray&& o) noexcept {
    if (this != &o) { Free(); p_ = o.p_; n_ = o.n_; o.p_ = nullptr; o.n_ = 0; }
    return *this;
  }
  ~DeviceArray() { Free(); }
  void Resize(size_t n) {
    Free();
    n_ = n;
    if (n > 0) PDLP_CUDA_CHECK(cudaMalloc(&p_, n * sizeof(T)));
  }
  void Free() {
    if (p_ != nullptr) cudaFree(p_);
    p_ = nullptr;
    n_ = 0;
  }
  T* get() const { return p_; }
  size_t size() const { return n_; }
  void Upload(const T* host, size_t n, cudaStream_t s) {
    if (n > 0) PDLP_CUDA_CHECK(cudaMemcpyAsync(p_, host, n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void Download(T* host, size_t n, cudaStream_t s) const {
    if (n > 0) PDLP_CUDA_CHECK(cudaMemcpyAsync(host, p_, n * sizeof(T), cudaMemcpyDeviceToHost, s));
  }
  void Zero(cudaStream_t s) {
    if (n_ > 0) PDLP_CUDA_CHECK(cudaMemsetAsync(p_, 0, n_ * sizeof(T), s));
  }
  void swap(DeviceArray& o) { std::swap(p_, o.p_); std::swap(n_, o.n_); }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};
using DVec = DeviceArray<double>;

constexpr int kBlock = 256;          // threads per CTA for every kernel here
constexpr int kMaxSlots = 16;        // reduction slots an op may use
constexpr int kStreamChunk = 2048;   // nonzeros staged in shared memory per work block
constexpr int kMaxGrid = 148 * 8;    // 148 SMs x 8 resident CTAs: persistent grid-stride launches

// Reduction slot kinds.  kMax slots must only ever receive values >= 0 or be
// initialised by the op through Op::Init (we initialise every slot with Identity()).
enum SlotKind : int { kSum = 0, kMax = 1 };

__device__ __forceinline__ double SlotIdentity(int kind) { return kind == kMax ? -INFINITY : 0.0; }
__device__ __forceinline__ double SlotCombine(int kind, double a, double b) {
  return kind == kMax ? fmax(a, b) : a + b;
}
// fmax drops NaNs; the stats pass detects NaNs through the sum slots, as the reference
// does through the l2 norm (primal_dual_hybrid_gradient.cc:737).

template <int NS>
struct Slots {
  double v[NS > 0 ? NS : 1];
};

// Block-wide reduction of NS slots; result valid in thread 0.  Fixed tree: lanes by
// xor-shuffle, then warps in index order.
template <class Op>
__device__ __forceinline__ void BlockReduceAndStore(Slots<Op::kNumSlots>& acc, double* partials) {
  constexpr int NS = Op::kNumSlots;
  if constexpr (NS > 0) {
    __shared__ double warp_buf[NS][kBlock / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      double x = acc.v[s];
      const int kind = Op::SlotKindOf(s);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) x = SlotCombine(kind, x, __shfl_xor_sync(0xffffffffu, x, off));
      if (lane == 0) warp_buf[s][warp] = x;
    }
    __syncthreads();
    if (threadIdx.x < NS) {
      const int s = threadIdx.x;
      const int kind = Op::SlotKindOf(s);
      double x = warp_buf[s][0];
      for (int w = 1; w < kBlock / 32; ++w) x = SlotCombine(kind, x, warp_buf[s][w]);
      partials[static_cast<size_t>(blockIdx.x) * NS + s] = x;
    }
  }
}

// ---- skeleton 1: dense index range ----
// Op interface:
//   static constexpr int kNumSlots; static __device__ int SlotKindOf(int s);
//   __device__ bool Begin();     // once per thread: resolve state-dependent pointers and
//                                // scalars into members; false = early-out (device-side
//                                // loop control)
//   __device__ void Apply(int64_t i, Slots<kNumSlots>& acc) const;
template <class Op>
__global__ void __launch_bounds__(kBlock) map_reduce_kernel(const Op op_in, int64_t count, double* partials) {
  Slots<Op::kNumSlots> acc;
#pragma unroll
  for (int s = 0; s < Op::kNumSlots; ++s) acc.v[s] = SlotIdentity(Op::SlotKindOf(s));
  Op op = op_in;
  if (op.Begin()) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kBlock;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * kBlock + threadIdx.x; i < count; i += stride) op.Apply(i, acc);
  }
  BlockReduceAndStore<Op>(acc, partials);
}

// ---- skeleton 2: compressed sparse matrix by rows ----
// The matrix is cut at setup into "work blocks" of consecutive rows holding at most
// kStreamChunk nonzeros (a row longer than that is a work block of its own).  A CTA
// streams its block's nonzeros with fully coalesced loads into shared memory as
// products Load(k), then sub-warps of T threads reduce each row out of shared memory
// (T = 1 for short rows: a sequential, reference-ordered sum).  Long rows are reduced
// straight from global memory by the whole CTA.
//
// Op interface (in addition to kNumSlots / SlotKindOf / Begin):
//   __device__ double Load(int64_t k) const;              // value staged for nonzero k
//   __device__ double RowInit() const;
//   __device__ double RowAdd(double acc, double staged, int row) const;
//   __device__ double RowMerge(double a, double b) const;
//   __device__ void RowFinish(int row, double acc, Slots<kNumSlots>& red) const;
struct RowBlocks {
  const int* row_ptr;      // rows + 1
  const int* block_start;  // num_blocks + 1 (row index where each work block starts)
  int num_blocks;
  int rows;
};

template <class Op>
__global__ void __launch_bounds__(kBlock) rowstream_kernel(const Op op_in, RowBlocks rb, double* partials) {
  __shared__ double staged[kStreamChunk];
  __shared__ double long_row_buf[kBlock / 32];
  Slots<Op::kNumSlots> red;
#pragma unroll
  for (int s = 0; s < Op::kNumSlots; ++s) red.v[s] = SlotIdentity(Op::SlotKindOf(s));
  Op op = op_in;
  if (op.Begin()) {
    for (int b = blockIdx.x; b < rb.num_blocks; b += gridDim.x) {
      const int r0 = rb.block_start[b], r1 = rb.block_start[b + 1];
      const int p0 = rb.row_ptr[r0], p1 = rb.row_ptr[r1];
      const int nz = p1 - p0;
      if (nz > kStreamChunk) {
        // One long row (r1 == r0 + 1 by construction): CTA-wide strided reduce.
        double a = op.RowInit();
        for (int k = p0 + threadIdx.x; k < p1; k += kBlock) a = op.RowAdd(a, op.Load(k), r0);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
        if ((threadIdx.x & 31) == 0) long_row_buf[threadIdx.x >> 5] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
          double t = long_row_buf[0];
          for (int w = 1; w < kBlock / 32; ++w) t = op.RowMerge(t, long_row_buf[w]);
          op.RowFinish(r0, t, red);
        }
        __syncthreads();
        continue;
      }
      // Stage: coalesced, 8 independent loads in flight per thread.
#pragma unroll 8
      for (int k = threadIdx.x; k < nz; k += kBlock) staged[k] = op.Load(p0 + k);
      __syncthreads();
      const int nrows = r1 - r0;
      // Threads per row: largest power of two <= min(32, kBlock / nrows), at least 1.
      int t = 1;
      while (t < 32 && t * 2 * nrows <= kBlock) t *= 2;
      const int sub = threadIdx.x / t, lane = threadIdx.x % t;
      const int groups = kBlock / t;
      for (int base = 0; base < nrows; base += groups) {
        const int r = r0 + base + sub;
        double a = op.RowInit();
        if (base + sub < nrows) {
          const int q0 = rb.row_ptr[r] - p0, q1 = rb.row_ptr[r + 1] - p0;
          for (int q = q0 + lane; q < q1; q += t) a = op.RowAdd(a, staged[q], r);
        }
        // All 32 lanes take part in the shuffles (t divides 32).
        for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
        if (lane == 0 && base + sub < nrows) op.RowFinish(r, a, red);
      }
      __syncthreads();
    }
  }
  BlockReduceAndStore<Op>(red, partials);
}

// ---- finalize: partials[num_blocks][NS] -> totals[NS], fixed order, one CTA ----
template <class Op>
__global__ void __launch_bounds__(kBlock) finalize_kernel(const double* partials, int num_blocks, double* totals) {
  constexpr int NS = Op::kNumSlots;
  __shared__ double buf[kBlock];
  for (int s = 0; s < NS; ++s) {
    const int kind = Op::SlotKindOf(s);
    double x = SlotIdentity(kind);
    for (int b = threadIdx.x; b < num_blocks; b += kBlock) x = SlotCombine(kind, x, partials[static_cast<size_t>(b) * NS + s]);
    buf[threadIdx.x] = x;
    __syncthreads();
    for (int off = kBlock / 2; off > 0; off >>= 1) {
      if (threadIdx.x < off) buf[threadIdx.x] = SlotCombine(kind, buf[threadIdx.x], buf[threadIdx.x + off]);
      __syncthreads();
    }
    if (threadIdx.x == 0) totals[s] = buf[0];
    __syncthreads();
  }
}

// Same reduction as a device function, for kernels that consume the totals themselves
// (the step-decision kernel).  Must be called by all kBlock threads of a single CTA.
__device__ __forceinline__ double ReducePartialsSum(const double* partials, int num_blocks, int stride, int slot, double* buf) {
  double x = 0.0;
  for (int b = threadIdx.x; b < num_blocks; b += kBlock) x += partials[static_cast<size_t>(b) * stride + slot];
  buf[threadIdx.x] = x;
  __syncthreads();
  for (int off = kBlock / 2; off > 0; off >>= 1) {
    if (threadIdx.x < off) buf[threadIdx.x] += buf[threadIdx.x + off];
    __syncthreads();
  }
  const double r = buf[0];
  __syncthreads();
  return r;
}

inline int GridFor(int64_t count) {
  int64_t g = (count + kBlock - 1) / kBlock;
  if (g < 1) g = 1;
  if (g > kMaxGrid) g = kMaxGrid;
  return static_cast<int>(g);
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_COMMON_CUH_
