// The row-stream pass of the hot ops (K x fused with the dual step, K^T y fused with the step-size
// test), software-pipelined over work blocks with asynchronous copies only.  Included by
// pdlp_skeletons.cuh; same work blocks, same Op epilogue interface and same per-row summation order
// as rowstream_kernel, which keeps serving the ops that compute their staged value with plain loads.
//
// Why a second skeleton.  rowstream_kernel brings a work block in with one burst of loads, waits for
// it, multiplies, synchronises and reduces: per CTA the memory phase and the reduction alternate, and
// ncu (profiles/r01_ncu_step_kernels_c3_v7.md) shows neither HBM (52-59 %), the issue slots (44-52 %) nor
// the L1 pipe saturated, 7-8 long-scoreboard stall cycles per issued instruction: the pass is bound by
// memory-level parallelism.  Cutting instructions alone (the "lean" staging of round 1: 79 -> 43 per
// nonzero) bought 6 % on K^T y and nothing on K x (profiles/r02_lean_ab.md).  Here every operand of
// block b + 1 - values, GATHERED vector entries, row offsets - and the indices of block b + 2 are in
// flight while block b is reduced:
//   * indices     global -> shared, 16-byte cp.async (four indices), two blocks ahead;
//   * values      global -> shared, 16-byte cp.async (two values), one block ahead;
//   * gather      x[idx] global -> shared, one 8-byte cp.async per nonzero with the index read from
//                 shared memory: no register holds a gathered value, so the copies of a whole block stay
//                 outstanding across the reduction of the previous one;
//   * row offsets global -> shared, 4-byte cp.async, one block ahead.
// The reduction multiplies value and gathered entry as it sums (no in-place product pass), one
// __syncthreads per work block.  128-bit copies need aligned windows: a block's window starts at its
// first nonzero rounded down to a multiple of four, so up to three nonzeros of the previous block are
// staged (and never summed) and up to three past the block's end are read: the matrix arrays carry four
// elements of padding (pdlp_matrix.cu), and a block holds at most kMaxStagedNnz = kStreamChunk - 4
// nonzeros so that its window fits kStreamChunk.
#ifndef PDLP_B200_ROWSTREAM_ASYNC_CUH_
#define PDLP_B200_ROWSTREAM_ASYNC_CUH_

namespace pdlp_b200 {

// Op interface of the asynchronous path (on top of kNumSlots / SlotKindOf / Begin / RowInit / RowAdd /
// RowMerge / RowFinish / optional Pre + RowPrefetch):
//   __device__ const int* IndexAddress(int64_t k) const;  __device__ const double* ValueAddress(int64_t k) const;
//   __device__ const double* GatherAddress(int index) const;   // &x[index]
//   __device__ double Product(double value, double gathered) const;
//   __device__ double Load(int64_t k) const;                   // rows longer than a work block
template <class Op, class = void>
struct HasAsyncGather : std::false_type {};
template <class Op>
struct HasAsyncGather<Op, std::void_t<decltype(std::declval<const Op&>().GatherAddress(0))>> : std::true_type {};

#ifndef PDLP_ASYNC_MIN_CTAS
#define PDLP_ASYNC_MIN_CTAS 2  // resident CTAs per SM the register allocation must allow (shared memory allows two at kStreamChunk = 2048)
#endif

struct AsyncSmem {  // dynamic shared memory of rowstream_async_kernel: two stages
  double val[2][kStagedSize];
  double gat[2][kStagedSize];
  int idx[2][kStreamChunk];
  int row_raw[2][kMaxRowsPerBlock + 4];
  double long_row_buf[kBlock / 32];
};

#ifdef PDLP_EMULATE
inline void CpAsync16(void* smem_dst, const void* global_src) { cuda_emu::AsyncCopy(smem_dst, global_src, 16); }
inline AsyncSmem* AsyncSmemBase() {
  static AsyncSmem s;  // one CTA at a time under the shim
  return &s;
}
#else
__device__ __forceinline__ void CpAsync16(void* smem_dst, const void* global_src) {
  const unsigned dst = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(global_src));
}
extern __shared__ __align__(16) unsigned char pdlp_async_smem_raw[];
__device__ __forceinline__ AsyncSmem* AsyncSmemBase() { return reinterpret_cast<AsyncSmem*>(pdlp_async_smem_raw); }
#endif

struct BlockMeta {
  int r0, r1, p0, p1;  // rows [r0, r1), nonzeros [p0, p1); all zero past the last work block
};
__device__ __forceinline__ BlockMeta LoadMeta(const RowBlocks& rb, int b) {
  BlockMeta m{0, 0, 0, 0};
  if (b < rb.num_blocks) { m.r0 = rb.block_start[b]; m.r1 = rb.block_start[b + 1]; m.p0 = rb.block_nz[b]; m.p1 = rb.block_nz[b + 1]; }
  return m;
}
// A block is staged through shared memory unless it is one row longer than a window.
__device__ __forceinline__ bool IsStaged(const BlockMeta& m) { return m.p1 - m.p0 > 0 && m.p1 - m.p0 <= kMaxStagedNnz; }

// Column indices of block m's window -> idx_s (four per copy).
template <class Op>
__device__ __forceinline__ void IssueIndices(const Op& op, int* idx_s, const BlockMeta& m) {
  if (!IsStaged(m)) return;
  const int a0 = m.p0 & ~3;
  const int nzw = m.p1 - a0;
  const int* src = op.IndexAddress(a0);
#pragma unroll
  for (int i = 0; i < kStreamChunk / (4 * kBlock); ++i) {
    const int e = 4 * (static_cast<int>(threadIdx.x) + i * kBlock);
    if (e < nzw) CpAsync16(idx_s + e, src + e);
  }
}

// Values, gathered entries and row offsets of block m -> stage `stage` (its indices are in idx_s already).
template <class Op>
__device__ __forceinline__ void IssueData(const Op& op, AsyncSmem& sm, int stage, const int* idx_s, const BlockMeta& m, const RowBlocks& rb) {
  const int nrows = m.r1 - m.r0;
  if (nrows == 0) return;
  int* row_s = sm.row_raw[stage];
#pragma unroll 1
  for (int i = threadIdx.x; i <= nrows; i += kBlock) CpAsync4(row_s + i, rb.row_ptr + m.r0 + i);
  if (!IsStaged(m)) return;
  const int a0 = m.p0 & ~3;
  const int nzw = m.p1 - a0;
  const int pad = BankPadFor(m.p1 - m.p0, nrows);
  double* val_s = sm.val[stage];
  double* gat_s = sm.gat[stage];
  const double* values = op.ValueAddress(a0);
  if (pad & 1) {
    // An odd pad moves every other group of 16 to an odd slot: pairs are only 8-byte aligned there.
#pragma unroll
    for (int i = 0; i < kPerThread; ++i) {
      const int e = static_cast<int>(threadIdx.x) + i * kBlock;
      if (e < nzw) CpAsync8(val_s + Skew(e, pad), values + e);
    }
  } else {
#pragma unroll
    for (int i = 0; i < kPerThread / 2; ++i) {
      const int e = 2 * (static_cast<int>(threadIdx.x) + i * kBlock);
      if (e < nzw) CpAsync16(val_s + Skew(e, pad), values + e);
    }
  }
  // Adjacent lanes take adjacent nonzeros, so runs of consecutive columns coalesce in the gather.
#pragma unroll
  for (int i = 0; i < kPerThread; ++i) {
    const int e = static_cast<int>(threadIdx.x) + i * kBlock;
    if (e < nzw) CpAsync8(gat_s + Skew(e, pad), op.GatherAddress(idx_s[e]));
  }
}

template <class Op>
__global__ void __launch_bounds__(kBlock, PDLP_ASYNC_MIN_CTAS) rowstream_async_kernel(const Op op_in, RowBlocks rb, double* partials) {
  AsyncSmem& sm = *AsyncSmemBase();
  using Pre = typename PreOf<Op>::type;
  constexpr int kPreRows = sizeof(Pre) > 16 ? 2 : 4;  // rows per thread whose epilogue operands are fetched before the reduction
  Slots<Op::kNumSlots> red;
#pragma unroll
  for (int s = 0; s < Op::kNumSlots; ++s) red.v[s] = SlotIdentity(Op::SlotKindOf(s));
  Op op = op_in;
  if (op.Begin()) {
    const int stride = gridDim.x;
    int b = blockIdx.x;
    BlockMeta cur = LoadMeta(rb, b), nxt = LoadMeta(rb, b + stride);
    // Prologue: this CTA's first block needs its indices before its gathers can be issued.
    IssueIndices(op, sm.idx[0], cur);
    CpAsyncCommit();
    CpAsyncWaitAll();
    __syncthreads();
    IssueData(op, sm, 0, sm.idx[0], cur, rb);
    IssueIndices(op, sm.idx[1], nxt);
    CpAsyncCommit();
    int s = 0;  // stage of the current block's data; its indices sat in idx[s], the next block's sit in idx[s ^ 1]
    while (b < rb.num_blocks) {
      const BlockMeta fut = LoadMeta(rb, b + 2 * stride);  // in flight for a whole block before it is needed
      // Everything issued an iteration ago has landed and is visible; every thread has left the previous
      // block's reduction, so stage s ^ 1 and idx[s] are free.
      CpAsyncWaitAll();
      __syncthreads();
      const int nz = cur.p1 - cur.p0, nrows = cur.r1 - cur.r0, r0 = cur.r0;
      // Threads per row: largest power of two <= min(32, kBlock / nrows), at least 1.
      int log_t = 0;
      while (log_t < 5 && (2 << log_t) * nrows <= kBlock) ++log_t;
      const int t = 1 << log_t;
      const int sub = threadIdx.x >> log_t, lane = threadIdx.x & (t - 1);
      const int groups = kBlock >> log_t;
      const bool staged = nz <= kMaxStagedNnz;
      // Epilogue operands first (their latency hides behind the issue phase and the reduction) ...
      Pre pre[kPreRows];
      if (staged) {
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (lane == 0 && rr < nrows) pre[u] = OpRowPrefetch(op, r0 + rr);
        }
      }
      // ... then the next block's operands and the indices of the one after it.
      IssueData(op, sm, s ^ 1, sm.idx[s ^ 1], nxt, rb);
      IssueIndices(op, sm.idx[s], fut);
      CpAsyncCommit();
      if (!staged) {
        // One long row (r1 == r0 + 1 by construction): CTA-wide strided reduce straight from global memory.
        const Pre lpre = OpRowPrefetch(op, r0);
        double a = op.RowInit();
        for (int k = cur.p0 + threadIdx.x; k < cur.p1; k += kBlock) a = op.RowAdd(a, op.Load(k), r0);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
        if ((threadIdx.x & 31) == 0) sm.long_row_buf[threadIdx.x >> 5] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
          double tot = sm.long_row_buf[0];
          for (int w = 1; w < kBlock / 32; ++w) tot = op.RowMerge(tot, sm.long_row_buf[w]);
          OpRowFinish(op, r0, tot, lpre, red);
        }
      } else {
        const int a0 = cur.p0 & ~3;
        const int pad = BankPadFor(nz > 0 ? nz : 1, nrows > 0 ? nrows : 1);
        const double* val_s = sm.val[s];
        const double* gat_s = sm.gat[s];
        const int* row_s = sm.row_raw[s];
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (u * groups < nrows) {  // uniform across the CTA
            double a = op.RowInit();
            if (rr < nrows) {
              const int q1 = row_s[rr + 1] - a0;
              for (int q = row_s[rr] - a0 + lane; q < q1; q += t) {
                const int pos = Skew(q, pad);
                a = op.RowAdd(a, op.Product(val_s[pos], gat_s[pos]), r0 + rr);
              }
            }
            // All 32 lanes take part in the shuffles (t divides 32).
            for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
            if (lane == 0 && rr < nrows) OpRowFinish(op, r0 + rr, a, pre[u], red);
          }
        }
        // Rows beyond the prefetched ones (only when a block has more than kPreRows * groups rows).
        for (int base = kPreRows * groups; base < nrows; base += groups) {
          const int rr = base + sub;
          double a = op.RowInit();
          if (rr < nrows) {
            const int q1 = row_s[rr + 1] - a0;
            for (int q = row_s[rr] - a0 + lane; q < q1; q += t) {
              const int pos = Skew(q, pad);
              a = op.RowAdd(a, op.Product(val_s[pos], gat_s[pos]), r0 + rr);
            }
          }
          for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
          if (lane == 0 && rr < nrows) OpRowFinish(op, r0 + rr, a, OpRowPrefetch(op, r0 + rr), red);
        }
      }
      b += stride;
      cur = nxt;
      nxt = fut;
      s ^= 1;
    }
    CpAsyncWaitAll();
  }
  BlockReduceAndStore<Op>(red, partials);
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_ROWSTREAM_ASYNC_CUH_
