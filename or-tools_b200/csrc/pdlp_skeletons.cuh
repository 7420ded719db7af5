// The two kernel skeletons every PDLP op is an instance of, and their deterministic block
// reductions (overview: pdlp_common.cuh).  Pure device code, no CUDA runtime calls: with
// PDLP_EMULATE defined (tests/kernel_emu/) the same source compiles as host C++ against a shim
// that runs a CTA as 256 host threads, so the index logic of the kernels is unit-tested on CPU.
#ifndef PDLP_B200_SKELETONS_CUH_
#define PDLP_B200_SKELETONS_CUH_

#ifndef PDLP_EMULATE
#include <cuda_runtime.h>
#endif

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <type_traits>
#include <utility>

namespace pdlp_b200 {

constexpr int kBlock = 256;          // threads per CTA for every kernel here
constexpr int kMaxSlots = 16;        // reduction slots an op may use
#ifndef PDLP_STREAM_CHUNK
#define PDLP_STREAM_CHUNK 2048
#endif
constexpr int kStreamChunk = PDLP_STREAM_CHUNK;   // nonzeros staged in shared memory per work block
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

// Base for ops that are always active and reduce nothing.
struct AlwaysActive {
  __device__ bool Begin() { return true; }
};
struct NoSlots {
  static constexpr int kNumSlots = 0;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
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
__global__ void __launch_bounds__(kBlock, 4) map_reduce_kernel(const Op op_in, int64_t count, double* partials) {
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
// products LoadAt(k, Index(k)), then sub-warps of T threads reduce each row out of shared
// memory (T = 1 for short rows: a sequential, reference-ordered sum).  Long rows are reduced
// straight from global memory by the whole CTA.
//
// The pass is latency-bound, not bandwidth-bound (profiles/r01_ncu_rowstream_c3_v0.md: DRAM at
// 51-57 % with full occupancy, stalls on the block_start -> row_ptr -> idx -> gather -> barrier
// -> row_ptr -> epilogue-operand chain of six dependent memory latencies per work block).  The
// kernel therefore software-pipelines everything that does not depend on the gather:
//   * work-block metadata (first row, first nonzero) two blocks ahead, from two flat arrays;
//   * the column indices of the NEXT block into shared memory (cp.async) while the current block
//     reduces, and the block's row offsets likewise;
//   * the epilogue operands of each thread's row(s) (Op::RowPrefetch) into registers, issued
//     BEFORE the gather so that the in-order warp does not stall in front of them;
// leaving idx (prefetched) -> gather -> barrier -> reduce from shared memory on the critical path.
// Shared-memory products are stored with a per-block bank pad so that the threads reducing rows of
// 8, 16, 32, 64 ... nonzeros do not collide on the same banks (see Skew below).
//
// Op interface (in addition to kNumSlots / SlotKindOf / Begin):
//   __device__ double Load(int64_t k) const;              // value staged for nonzero k
//   __device__ double RowInit() const;
//   __device__ double RowAdd(double acc, double staged, int row) const;
//   __device__ double RowMerge(double a, double b) const;
//   __device__ void RowFinish(int row, double acc, Slots<kNumSlots>& red) const;
// Optional (the hot ops have them):
//   __device__ const int* IndexAddress(int64_t k) const;  __device__ const double* ValueAddress(int64_t k) const;
//   __device__ double Gather(int index) const;  __device__ double Product(double value, double gathered) const;
//       split of Load (= Product(*ValueAddress(k), Gather(*IndexAddress(k)))) so that indices travel one
//       block ahead and values by asynchronous copy
//   struct Pre;  __device__ Pre RowPrefetch(int row) const;
//   __device__ void RowFinish(int row, double acc, const Pre& pre, Slots<kNumSlots>& red) const;
//       epilogue operands fetched before the reduction instead of after it
struct RowBlocks {
  const int* row_ptr;      // rows + 1
  const int* block_start;  // num_blocks + 1 (row index where each work block starts)
  const int* block_nz;     // num_blocks + 1 (row_ptr[block_start[b]]: first nonzero of each work block)
  int num_blocks;
  int rows;
  // Ops that push row sums to other GPUs (HasBlockPush): CTAs [0, split_ctas) take the work blocks [0, split_block) - the
  // rows whose sums travel over NVLink - and the other CTAs the rest, AT THE SAME TIME, so the link is busy for the whole
  // launch instead of a burst at its start that every CTA then waits behind.  split_ctas = 0: one sequence over all CTAs.
  int split_block = 0;
  int split_ctas = 0;
};

constexpr int kMaxRowsPerBlock = 4 * kBlock;
#ifndef PDLP_ROWSTREAM_MIN_CTAS
#define PDLP_ROWSTREAM_MIN_CTAS 4  // resident CTAs per SM the register allocation must allow
#endif
constexpr int kPerThread = kStreamChunk / kBlock;             // nonzeros per thread per work block
constexpr int kStagedSize = kStreamChunk + 3 * (kStreamChunk / 16);  // with the largest bank pad
// Position of staged product k: c pad slots after every 16 products, c = 0 .. 3 chosen per work
// block.  With 8-byte products a half-warp touches 16 bank pairs.  The three coalesced passes over
// the products (asynchronous copy in, multiply in place) walk 16 consecutive k that start at a multiple of 16, so a
// pad that only falls between such groups keeps them conflict-free (a pad every 8 products does
// not: measured +6 % on K x).  The row reduction reads T threads per row; simulated conflict
// degrees (scripts/bank_pad_sim.py, wavefronts per 16 products, 1.0 = none):
//   rows of one odd length (3, 5: CSC copies of the bench LPs)    c=0: 1.0   c=1: 3-5   c=2,3: 2.0
//   rows of one length that is a multiple of 16 (16, 32, 64)       c=0: 2-8   c=1: 2.0   c=2: 1.0   c=3: 2.0
//   rows of one other even length (2, 4, 8, 10, 12, 24)            c=0: 2-8   c=1: 1-5   c=2: 2-3   c=3: 1.0-2.0
//   irregular rows (Poisson lengths around 10 or 16)               c=0: 3.6   c=1: 3.5   c=2: 3.1-3.6   c=3: 3.0-3.6
// A block is "of one length" when nz == nrows * (nz / nrows).  (Before this rule, c = 1 everywhere:
// 24 % of the shared-memory wavefronts of both SpMV kernels on C3 were bank conflicts,
// profiles/r01_ncu_step_kernels_c3_v6.md.)
__device__ __forceinline__ int Skew(int k, int pad) { return k + pad * (k >> 4); }
__device__ __forceinline__ int BankPadFor(int nz, int nrows) {
  const int len = nz / nrows;
  if (len * nrows != nz) return 3;
  if (len & 1) return 0;
  return (len & 15) == 0 ? 2 : 3;
}

template <class Op, class = void>
struct HasSplitLoad : std::false_type {};
template <class Op>
struct HasSplitLoad<Op, std::void_t<decltype(std::declval<const Op&>().IndexAddress(int64_t{0}))>> : std::true_type {};
template <class Op, class = void>
struct HasRowPrefetch : std::false_type {};
template <class Op>
struct HasRowPrefetch<Op, std::void_t<typename Op::Pre>> : std::true_type {};

// An op that hands the sums of some rows to other GPUs (pdlp_peer.cuh) declares `bool PushesRow(int row)` and
// `void BlockPush(int r0, int nrows, const double* sums)`: the kernel then collects those rows' sums of a work block in
// shared memory and the whole CTA stores them, so that a warp writes 256 contiguous bytes per destination instead of
// one 8-byte store per row from a single lane (NVLink moves packets, and 8-byte packets are mostly header).
template <class Op, class = void>
struct HasBlockPush : std::false_type {};
template <class Op>
struct HasBlockPush<Op, std::void_t<decltype(std::declval<const Op&>().PushesRow(0))>> : std::true_type {};

struct NoPre {};
template <class Op, bool = HasRowPrefetch<Op>::value>
struct PreOf { using type = NoPre; };
template <class Op>
struct PreOf<Op, true> { using type = typename Op::Pre; };

template <class Op>
__device__ __forceinline__ typename PreOf<Op>::type OpRowPrefetch(const Op& op, int row) {
  if constexpr (HasRowPrefetch<Op>::value) return op.RowPrefetch(row);
  else return NoPre{};
}
template <class Op>
__device__ __forceinline__ void OpRowFinish(const Op& op, int row, double acc, const typename PreOf<Op>::type& pre, Slots<Op::kNumSlots>& red) {
  if constexpr (HasRowPrefetch<Op>::value) op.RowFinish(row, acc, pre, red);
  else op.RowFinish(row, acc, red);
}

// Ampere-style asynchronous 4-byte copies global -> shared: the row offsets of the current work
// block and the column indices of the next one travel without occupying registers or stalling
// the issuing warp.
#ifdef PDLP_EMULATE
inline void CpAsync4(void* smem_dst, const void* global_src) { cuda_emu::AsyncCopy(smem_dst, global_src, 4); }
inline void CpAsync8(void* smem_dst, const void* global_src) { cuda_emu::AsyncCopy(smem_dst, global_src, 8); }
inline void CpAsyncCommit() {}
inline void CpAsyncWaitAll() { cuda_emu::AsyncWaitAll(); }
#else
__device__ __forceinline__ void CpAsync4(void* smem_dst, const void* global_src) {
  const unsigned dst = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(global_src));
}
__device__ __forceinline__ void CpAsync8(void* smem_dst, const void* global_src) {
  const unsigned dst = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(dst), "l"(global_src));
}
__device__ __forceinline__ void CpAsyncCommit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void CpAsyncWaitAll() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }
#endif

template <class Op>
__device__ __forceinline__ void PrefetchIndices(const Op& op, int* idx_buf, int first_nz, int nz) {
  if constexpr (HasSplitLoad<Op>::value) {
    if (nz <= kStreamChunk) {
      {
#pragma unroll
        for (int i = 0; i < kPerThread; ++i) {
          const int k = threadIdx.x + i * kBlock;
          if (k < nz) CpAsync4(idx_buf + k, op.IndexAddress(first_nz + k));
        }
      }
    }
  }
  CpAsyncCommit();
}

template <class Op>
__global__ void __launch_bounds__(kBlock, PDLP_ROWSTREAM_MIN_CTAS) rowstream_kernel(const Op op_in, RowBlocks rb, double* partials) {
  __shared__ double staged[kStagedSize];
  __shared__ int row_raw[kMaxRowsPerBlock + 1];  // row_ptr[r0 ...] of the current work block
  // One buffer is enough for "this block's indices" and "the next block's indices": a thread copies
  // the next block's indices into exactly the positions it has already read this block's from.
  __shared__ int idx_buf[HasSplitLoad<Op>::value ? kStreamChunk : 1];
  __shared__ double long_row_buf[kBlock / 32];
  __shared__ double push_buf[HasBlockPush<Op>::value ? kMaxRowsPerBlock : 1];  // sums of this block's pushed rows
  using Pre = typename PreOf<Op>::type;
  constexpr int kPreRows = sizeof(Pre) > 16 ? 2 : 4;  // rows per thread whose epilogue operands are fetched before the barrier
  Slots<Op::kNumSlots> red;
#pragma unroll
  for (int s = 0; s < Op::kNumSlots; ++s) red.v[s] = SlotIdentity(Op::SlotKindOf(s));
  Op op = op_in;
  if (op.Begin()) {
    int stride = gridDim.x;
    int b = blockIdx.x;
    int end = rb.num_blocks;
    if constexpr (HasBlockPush<Op>::value) {
      if (rb.split_ctas > 0) {
        if (static_cast<int>(blockIdx.x) < rb.split_ctas) { stride = rb.split_ctas; end = rb.split_block; }
        else { b = rb.split_block + (blockIdx.x - rb.split_ctas); stride = gridDim.x - rb.split_ctas; }
      }
    }
    // Metadata of the current work block and of the next one (b + stride).
    int r0 = 0, r1 = 0, p0 = 0, p1 = 0, nr0 = 0, nr1 = 0, np0 = 0, np1 = 0;
    if (b < end) { r0 = rb.block_start[b]; r1 = rb.block_start[b + 1]; p0 = rb.block_nz[b]; p1 = rb.block_nz[b + 1]; }
    if (b + stride < end) {
      nr0 = rb.block_start[b + stride]; nr1 = rb.block_start[b + stride + 1];
      np0 = rb.block_nz[b + stride]; np1 = rb.block_nz[b + stride + 1];
    }
    PrefetchIndices(op, idx_buf, p0, p1 - p0);
    while (b < end) {
      const int nz = p1 - p0;
      // Metadata two blocks ahead: in flight for a whole block before it is needed.
      const int b2 = b + 2 * stride;
      int fr0 = 0, fr1 = 0, fp0 = 0, fp1 = 0;
      if (b2 < end) { fr0 = rb.block_start[b2]; fr1 = rb.block_start[b2 + 1]; fp0 = rb.block_nz[b2]; fp1 = rb.block_nz[b2 + 1]; }
      const int next_nz = np1 - np0;  // 0 when there is no next block
      if (nz > kStreamChunk) {
        // One long row (r1 == r0 + 1 by construction): CTA-wide strided reduce.
        PrefetchIndices(op, idx_buf, np0, next_nz);
        const Pre pre = OpRowPrefetch(op, r0);
        double a = op.RowInit();
        for (int k = p0 + threadIdx.x; k < p1; k += kBlock) a = op.RowAdd(a, op.Load(k), r0);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
        if ((threadIdx.x & 31) == 0) long_row_buf[threadIdx.x >> 5] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
          double t = long_row_buf[0];
          for (int w = 1; w < kBlock / 32; ++w) t = op.RowMerge(t, long_row_buf[w]);
          OpRowFinish(op, r0, t, pre, red);
        }
        CpAsyncWaitAll();
        __syncthreads();
      } else {
        // This block's column indices were copied to shared memory a block ago (own positions).
        CpAsyncWaitAll();
        int jx[kPerThread];
        if constexpr (HasSplitLoad<Op>::value) {
#pragma unroll
          for (int i = 0; i < kPerThread; ++i) {
            const int k = threadIdx.x + i * kBlock;
            jx[i] = k < nz ? idx_buf[k] : 0;
          }
        }
        // Issue order matters (warps issue in order, and ptxas places the cp.async wait directly
        // behind the last copy): epilogue operands and gathers first - their addresses are known -
        // then the asynchronous copies of this block's values and row offsets and of the next
        // block's indices.  All of them are in flight together; the wait then covers one latency.
        const int nrows = r1 - r0;
        const int shift = BankPadFor(nz, nrows);
        // Threads per row: largest power of two <= min(32, kBlock / nrows), at least 1.
        int log_t = 0;
        while (log_t < 5 && (2 << log_t) * nrows <= kBlock) ++log_t;
        const int t = 1 << log_t;
        const int sub = threadIdx.x >> log_t, lane = threadIdx.x & (t - 1);
        const int groups = kBlock >> log_t;
        Pre pre[kPreRows];
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (lane == 0 && rr < nrows) pre[u] = OpRowPrefetch(op, r0 + rr);
        }
        if constexpr (HasSplitLoad<Op>::value) {
          double g[kPerThread];
#pragma unroll
          for (int i = 0; i < kPerThread; ++i) {
            const int k = threadIdx.x + i * kBlock;
            g[i] = k < nz ? op.Gather(jx[i]) : 0.0;
          }
#pragma unroll
          for (int i = 0; i < kPerThread; ++i) {
            const int k = threadIdx.x + i * kBlock;
            if (k < nz) CpAsync8(staged + Skew(k, shift), op.ValueAddress(p0 + k));
          }
          for (int i = threadIdx.x; i <= nrows; i += kBlock) CpAsync4(row_raw + i, rb.row_ptr + r0 + i);
          PrefetchIndices(op, idx_buf, np0, next_nz);
          CpAsyncWaitAll();
          // Each thread multiplies its own positions in place.
#pragma unroll
          for (int i = 0; i < kPerThread; ++i) {
            const int k = threadIdx.x + i * kBlock;
            if (k < nz) staged[Skew(k, shift)] = op.Product(staged[Skew(k, shift)], g[i]);
          }
        } else {
          for (int i = threadIdx.x; i <= nrows; i += kBlock) CpAsync4(row_raw + i, rb.row_ptr + r0 + i);
          CpAsyncCommit();
#pragma unroll
          for (int i = 0; i < kPerThread; ++i) {
            const int k = threadIdx.x + i * kBlock;
            if (k < nz) staged[Skew(k, shift)] = op.Load(p0 + k);
          }
          CpAsyncWaitAll();
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (u * groups < nrows) {  // uniform across the CTA
            double a = op.RowInit();
            if (rr < nrows) {
              const int q1 = row_raw[rr + 1] - p0;
              for (int q = row_raw[rr] - p0 + lane; q < q1; q += t) a = op.RowAdd(a, staged[Skew(q, shift)], r0 + rr);
            }
            // All 32 lanes take part in the shuffles (t divides 32).
            for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
            if (lane == 0 && rr < nrows) {
              if constexpr (HasBlockPush<Op>::value) {
                if (op.PushesRow(r0 + rr)) push_buf[rr] = a;
                else OpRowFinish(op, r0 + rr, a, pre[u], red);
              } else {
                OpRowFinish(op, r0 + rr, a, pre[u], red);
              }
            }
          }
        }
        // Rows beyond the prefetched ones (only when a block has more than kPreRows * groups rows).
        for (int base = kPreRows * groups; base < nrows; base += groups) {
          const int rr = base + sub;
          double a = op.RowInit();
          if (rr < nrows) {
            const int q1 = row_raw[rr + 1] - p0;
            for (int q = row_raw[rr] - p0 + lane; q < q1; q += t) a = op.RowAdd(a, staged[Skew(q, shift)], r0 + rr);
          }
          for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
          if (lane == 0 && rr < nrows) {
            if constexpr (HasBlockPush<Op>::value) {
              if (op.PushesRow(r0 + rr)) push_buf[rr] = a;
              else OpRowFinish(op, r0 + rr, a, OpRowPrefetch(op, r0 + rr), red);
            } else {
              OpRowFinish(op, r0 + rr, a, OpRowPrefetch(op, r0 + rr), red);
            }
          }
        }
        __syncthreads();
        // A work block lies on one side of the shared / private cut.  The next block writes push_buf only behind its
        // own barrier, which no thread passes before every thread has left this loop.
        if constexpr (HasBlockPush<Op>::value) {
          if (op.PushesRow(r0)) op.BlockPush(r0, nrows, push_buf);
        }
      }
      b += stride;
      r0 = nr0; r1 = nr1; p0 = np0; p1 = np1;
      nr0 = fr0; nr1 = fr1; np0 = fp0; np1 = fp1;
    }
    CpAsyncWaitAll();
  }
  if constexpr (HasBlockPush<Op>::value) {
    // A split launch signals from inside (pdlp_peer.cuh:PeerSharedCtaDone), switched-off op included.
    if (rb.split_ctas > 0 && static_cast<int>(blockIdx.x) < rb.split_ctas) op.PushingCtaDone(rb.split_ctas);
  }
  BlockReduceAndStore<Op>(red, partials);
}

// ---- skeleton 2b: several products with ONE stream of the matrix ----
// out_p[row] = sum_k val[k] v_p[idx[k]] for p = 0 .. kPoints - 1.  A thread keeps its nonzeros' values and indices in
// registers and the work block is staged, reduced and stored once per point, so the matrix (12 bytes per nonzero, the
// bulk of a pass) is read once and only the gathers - served by L1 / L2 - repeat.  Used at host events, where the
// termination statistics and the restart bounds need K x and K^T y of the current iterate, the average and the last
// difference at the same time (pdlp_kkt.cu).  Every row sum is formed exactly as rowstream_kernel forms it (same
// threads per row, same order of additions, same shuffle tree), so the sums are bit-identical to single passes.
// Op interface: kPoints; Begin(); double Value(int64_t k); int Index(int64_t k); double Gather(int point, int index);
//               void Store(int point, int row, double sum).
template <class Op>
__global__ void __launch_bounds__(kBlock, PDLP_ROWSTREAM_MIN_CTAS) rowstream_multi_kernel(const Op op_in, RowBlocks rb) {
  __shared__ double staged[kStagedSize];
  __shared__ int row_raw[kMaxRowsPerBlock + 1];
  __shared__ double long_row_buf[kBlock / 32];
  Op op = op_in;
  if (!op.Begin()) return;
  for (int b = blockIdx.x; b < rb.num_blocks; b += gridDim.x) {
    const int r0 = rb.block_start[b], r1 = rb.block_start[b + 1];
    const int p0 = rb.block_nz[b], p1 = rb.block_nz[b + 1];
    const int nz = p1 - p0, nrows = r1 - r0;
    if (nz > kStreamChunk) {
      // One long row: CTA-wide strided reduce per point (the row's values come from L2 after the first point).
#pragma unroll
      for (int pt = 0; pt < Op::kPoints; ++pt) {
        double a = 0.0;
        for (int k = p0 + threadIdx.x; k < p1; k += kBlock) a = a + op.Value(k) * op.Gather(pt, op.Index(k));
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a = a + __shfl_xor_sync(0xffffffffu, a, off);
        if ((threadIdx.x & 31) == 0) long_row_buf[threadIdx.x >> 5] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
          double t = long_row_buf[0];
          for (int w = 1; w < kBlock / 32; ++w) t = t + long_row_buf[w];
          op.Store(pt, r0, t);
        }
        __syncthreads();
      }
      continue;
    }
    double v[kPerThread];
    int jx[kPerThread];
#pragma unroll
    for (int i = 0; i < kPerThread; ++i) {
      const int k = threadIdx.x + i * kBlock;
      v[i] = k < nz ? op.Value(p0 + k) : 0.0;
      jx[i] = k < nz ? op.Index(p0 + k) : 0;
    }
    for (int i = threadIdx.x; i <= nrows; i += kBlock) row_raw[i] = rb.row_ptr[r0 + i];
    const int shift = BankPadFor(nz, nrows);
    int log_t = 0;
    while (log_t < 5 && (2 << log_t) * nrows <= kBlock) ++log_t;
    const int t = 1 << log_t;
    const int sub = threadIdx.x >> log_t, lane = threadIdx.x & (t - 1);
    const int groups = kBlock >> log_t;
#pragma unroll
    for (int pt = 0; pt < Op::kPoints; ++pt) {
#pragma unroll
      for (int i = 0; i < kPerThread; ++i) {
        const int k = threadIdx.x + i * kBlock;
        if (k < nz) staged[Skew(k, shift)] = v[i] * op.Gather(pt, jx[i]);
      }
      __syncthreads();
      for (int base = 0; base < nrows; base += groups) {
        const int rr = base + sub;
        double a = 0.0;
        if (rr < nrows) {
          const int q1 = row_raw[rr + 1] - p0;
          for (int q = row_raw[rr] - p0 + lane; q < q1; q += t) a = a + staged[Skew(q, shift)];
        }
        for (int off = t >> 1; off > 0; off >>= 1) a = a + __shfl_xor_sync(0xffffffffu, a, off);
        if (lane == 0 && rr < nrows) op.Store(pt, r0 + rr, a);
      }
      __syncthreads();
    }
  }
}

// ---- the two phases of a column-sharded pass over the CSR block (launched by
// pdlp_launch.cuh:EnqueueRows around the all-reduce of the shared rows' sums) ----
// An op whose row accumulator combines with max instead of + declares `bool RowCombineMax() const`.
template <class Op, class = void>
struct HasRowCombineMax : std::false_type {};
template <class Op>
struct HasRowCombineMax<Op, std::void_t<decltype(std::declval<const Op&>().RowCombineMax())>> : std::true_type {};
template <class Op>
inline bool RowCombineIsMax(const Op& op) {
  if constexpr (HasRowCombineMax<Op>::value) return op.RowCombineMax();
  else return false;
}

// Phase 1 of a sharded CSR pass: this rank's part of every row sum, stored.
template <class Op>
struct PartialRowOp {
  static constexpr int kNumSlots = 0;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  Op op;
  double* out;
  __device__ bool Begin() { return op.Begin(); }
  __device__ double Load(int64_t k) const { return op.Load(k); }
  template <class O = Op, class = std::enable_if_t<HasSplitLoad<O>::value>>
  __device__ const int* IndexAddress(int64_t k) const { return op.IndexAddress(k); }
  template <class O = Op, class = std::enable_if_t<HasSplitLoad<O>::value>>
  __device__ const double* ValueAddress(int64_t k) const { return op.ValueAddress(k); }
  template <class O = Op, class = std::enable_if_t<HasSplitLoad<O>::value>>
  __device__ double Gather(int index) const { return op.Gather(index); }
  template <class O = Op, class = std::enable_if_t<HasSplitLoad<O>::value>>
  __device__ double Product(double value, double gathered) const { return op.Product(value, gathered); }
  __device__ double RowInit() const { return op.RowInit(); }
  __device__ double RowAdd(double acc, double staged, int row) const { return op.RowAdd(acc, staged, row); }
  __device__ double RowMerge(double a, double b) const { return op.RowMerge(a, b); }
  __device__ void RowFinish(int row, double acc, Slots<0>&) const { out[row] = acc; }
};
// Phase 2: the op's epilogue on the all-reduced row sums.
template <class Op>
struct FinishRowOp {
  static constexpr int kNumSlots = Op::kNumSlots;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return Op::SlotKindOf(s); }
  Op op;
  const double* sums;
  __device__ bool Begin() { return op.Begin(); }
  __device__ void Apply(int64_t i, Slots<Op::kNumSlots>& acc) const {
    const int row = static_cast<int>(i);
    OpRowFinish(op, row, sums[i], OpRowPrefetch(op, row), acc);
  }
};

}  // namespace pdlp_b200
#include "pdlp_peer.cuh"  // the same two phases over peer memory instead of an all-reduce
namespace pdlp_b200 {

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
#endif  // PDLP_B200_SKELETONS_CUH_
