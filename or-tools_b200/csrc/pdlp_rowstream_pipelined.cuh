// Register-pipelined row-stream pass for the split-load ops (K x with the dual step, K^T y with the
// step-size test).  Included by pdlp_skeletons.cuh.  Same work blocks, staged layout, per-row summation
// order and Op interface as rowstream_kernel's split-load path; what changes is WHEN the loads are issued:
//
//   rowstream_kernel, block b:   gathers(b) + values(b) -> wait -> products -> barrier -> reduce(b) -> barrier
//   this kernel, block b:        wait -> products(b) -> barrier -> gathers(b+1) + values(b+1) + indices(b+2)
//                                issued -> reduce(b)
//
// so every CTA has a block's worth of loads in flight while it reduces, and a block costs one barrier.
// The gathered entries of block b + 1 wait in registers (kPerThread doubles per thread), the values in
// the second half of a double-buffered staging array.  Gathers are plain loads on purpose: per-element
// cp.async gathers were measured 1.3 - 2.6 x slower (profiles/r02_rowstream_ab.md).
#ifndef PDLP_B200_ROWSTREAM_PIPELINED_CUH_
#define PDLP_B200_ROWSTREAM_PIPELINED_CUH_

namespace pdlp_b200 {

#ifndef PDLP_PIPE_MIN_CTAS
#define PDLP_PIPE_MIN_CTAS 3  // resident CTAs per SM the register allocation must allow
#endif

struct PipeSmem {  // dynamic shared memory of rowstream_pipelined_kernel (55 KB: above the static limit)
  double staged[2][kStagedSize];
  int row_raw[2][kMaxRowsPerBlock + 1];
  int idx_buf[kStreamChunk];  // one buffer: a thread overwrites exactly the positions it has read
  double long_row_buf[kBlock / 32];
};
#ifdef PDLP_EMULATE
inline PipeSmem* PipeSmemBase() {
  static PipeSmem s;  // one CTA at a time under the shim
  return &s;
}
#else
extern __shared__ __align__(16) unsigned char pdlp_pipe_smem_raw[];
__device__ __forceinline__ PipeSmem* PipeSmemBase() { return reinterpret_cast<PipeSmem*>(pdlp_pipe_smem_raw); }
#endif

struct PipeMeta {
  int r0, r1, p0, p1;  // rows [r0, r1), nonzeros [p0, p1); all zero past the last work block
};
__device__ __forceinline__ PipeMeta PipeLoadMeta(const RowBlocks& rb, int b) {
  PipeMeta m{0, 0, 0, 0};
  if (b < rb.num_blocks) { m.r0 = rb.block_start[b]; m.r1 = rb.block_start[b + 1]; m.p0 = rb.block_nz[b]; m.p1 = rb.block_nz[b + 1]; }
  return m;
}
// A block goes through shared memory unless it is one row longer than the staging array (or the empty
// block past the end of the matrix).
__device__ __forceinline__ bool PipeStaged(const PipeMeta& m) { return m.r1 > m.r0 && m.p1 - m.p0 <= kStreamChunk; }

// Issues everything block m needs (its indices are in idx_buf at the thread's own positions): gathers into
// g[], asynchronous copies of the values into `staged` and of the row offsets into `row_raw`.  Returns the
// block's bank pad.
template <class Op>
__device__ __forceinline__ int PipeIssue(const Op& op, const PipeMeta& m, const RowBlocks& rb, const int* idx_buf, double* staged, int* row_raw,
                                         double (&g)[kPerThread]) {
  const int nz = m.p1 - m.p0, nrows = m.r1 - m.r0;
  if (!PipeStaged(m)) return 0;
  const int pad = BankPadFor(nz, nrows);
#pragma unroll
  for (int i = 0; i < kPerThread; ++i) {
    const int k = threadIdx.x + i * kBlock;
    g[i] = k < nz ? op.Gather(idx_buf[k]) : 0.0;
  }
#pragma unroll
  for (int i = 0; i < kPerThread; ++i) {
    const int k = threadIdx.x + i * kBlock;
    if (k < nz) CpAsync8(staged + Skew(k, pad), op.ValueAddress(m.p0 + k));
  }
  for (int i = threadIdx.x; i <= nrows; i += kBlock) CpAsync4(row_raw + i, rb.row_ptr + m.r0 + i);
  return pad;
}

template <class Op>
__global__ void __launch_bounds__(kBlock, PDLP_PIPE_MIN_CTAS) rowstream_pipelined_kernel(const Op op_in, RowBlocks rb, double* partials) {
  PipeSmem& sm = *PipeSmemBase();
  double (*staged)[kStagedSize] = sm.staged;
  int (*row_raw)[kMaxRowsPerBlock + 1] = sm.row_raw;
  int* idx_buf = sm.idx_buf;
  double* long_row_buf = sm.long_row_buf;
  using Pre = typename PreOf<Op>::type;
  constexpr int kPreRows = sizeof(Pre) > 16 ? 2 : 4;
  Slots<Op::kNumSlots> red;
#pragma unroll
  for (int s = 0; s < Op::kNumSlots; ++s) red.v[s] = SlotIdentity(Op::SlotKindOf(s));
  Op op = op_in;
  if (op.Begin()) {
    const int stride = gridDim.x;
    int b = blockIdx.x;
    PipeMeta cur = PipeLoadMeta(rb, b), nxt = PipeLoadMeta(rb, b + stride);
    double g[kPerThread];
    // Prologue: the first block's indices, then its loads and the second block's indices.
    PrefetchIndices(op, idx_buf, cur.p0, PipeStaged(cur) ? cur.p1 - cur.p0 : 0);
    CpAsyncWaitAll();
    int pad = PipeIssue(op, cur, rb, idx_buf, staged[0], row_raw[0], g);
    PrefetchIndices(op, idx_buf, nxt.p0, PipeStaged(nxt) ? nxt.p1 - nxt.p0 : 0);
    int s = 0;
    while (b < rb.num_blocks) {
      const PipeMeta fut = PipeLoadMeta(rb, b + 2 * stride);
      const int nz = cur.p1 - cur.p0, nrows = cur.r1 - cur.r0, r0 = cur.r0;
      const bool is_staged = PipeStaged(cur);
      // This block's values and offsets and the next block's indices have landed (own copies).
      CpAsyncWaitAll();
      if (is_staged) {
#pragma unroll
        for (int i = 0; i < kPerThread; ++i) {
          const int k = threadIdx.x + i * kBlock;
          if (k < nz) staged[s][Skew(k, pad)] = op.Product(staged[s][Skew(k, pad)], g[i]);
        }
      }
      // Products and offsets of this block visible to all; every thread has left the previous block's
      // reduction, so the other staging half is free.
      __syncthreads();
      int log_t = 0;
      while (log_t < 5 && (2 << log_t) * nrows <= kBlock) ++log_t;
      const int t = 1 << log_t;
      const int sub = threadIdx.x >> log_t, lane = threadIdx.x & (t - 1);
      const int groups = kBlock >> log_t;
      Pre pre[kPreRows];
      if (is_staged) {
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (lane == 0 && rr < nrows) pre[u] = OpRowPrefetch(op, r0 + rr);
        }
      }
      // The next block's loads go out before this block is reduced.
      const int next_pad = PipeIssue(op, nxt, rb, idx_buf, staged[s ^ 1], row_raw[s ^ 1], g);
#ifndef PDLP_EMULATE
      asm volatile("" ::: "memory");  // the index reads above stay ahead of the copies that overwrite idx_buf
#endif
      PrefetchIndices(op, idx_buf, fut.p0, PipeStaged(fut) ? fut.p1 - fut.p0 : 0);
      if (!is_staged) {
        if (nrows > 0) {
          // One long row (r1 == r0 + 1 by construction): CTA-wide strided reduce straight from global memory.
          // long_row_buf is free: thread 0 read it before it arrived at this iteration's barrier.
          const Pre lpre = OpRowPrefetch(op, r0);
          double a = op.RowInit();
          for (int k = cur.p0 + threadIdx.x; k < cur.p1; k += kBlock) a = op.RowAdd(a, op.Load(k), r0);
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
          if ((threadIdx.x & 31) == 0) long_row_buf[threadIdx.x >> 5] = a;
          __syncthreads();
          if (threadIdx.x == 0) {
            double tot = long_row_buf[0];
            for (int w = 1; w < kBlock / 32; ++w) tot = op.RowMerge(tot, long_row_buf[w]);
            OpRowFinish(op, r0, tot, lpre, red);
          }
        }
      } else {
        const double* stg = staged[s];
        const int* rows_s = row_raw[s];
#pragma unroll
        for (int u = 0; u < kPreRows; ++u) {
          const int rr = sub + u * groups;
          if (u * groups < nrows) {  // uniform across the CTA
            double a = op.RowInit();
            if (rr < nrows) {
              const int q1 = rows_s[rr + 1] - cur.p0;
              for (int q = rows_s[rr] - cur.p0 + lane; q < q1; q += t) a = op.RowAdd(a, stg[Skew(q, pad)], r0 + rr);
            }
            for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
            if (lane == 0 && rr < nrows) OpRowFinish(op, r0 + rr, a, pre[u], red);
          }
        }
        for (int base = kPreRows * groups; base < nrows; base += groups) {
          const int rr = base + sub;
          double a = op.RowInit();
          if (rr < nrows) {
            const int q1 = rows_s[rr + 1] - cur.p0;
            for (int q = rows_s[rr] - cur.p0 + lane; q < q1; q += t) a = op.RowAdd(a, stg[Skew(q, pad)], r0 + rr);
          }
          for (int off = t >> 1; off > 0; off >>= 1) a = op.RowMerge(a, __shfl_xor_sync(0xffffffffu, a, off));
          if (lane == 0 && rr < nrows) OpRowFinish(op, r0 + rr, a, OpRowPrefetch(op, r0 + rr), red);
        }
      }
      b += stride;
      cur = nxt;
      nxt = fut;
      pad = next_pad;
      s ^= 1;
    }
    CpAsyncWaitAll();
  }
  BlockReduceAndStore<Op>(red, partials);
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_ROWSTREAM_PIPELINED_CUH_
