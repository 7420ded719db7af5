// Peer-memory exchange of the column-sharded solve: the replacement of the two
// ncclAllReduce calls inside a trial step.  Device side; the host side (allocation, CUDA IPC handles,
// epochs) is PeerExchange in pdlp_comm.h.
//
// Every rank owns one region in its own HBM that all its peers have mapped (one process per GPU, CUDA
// IPC).  A sender PUSHES: its row-stream kernel stores each partial row sum into slot [its rank] of every
// rank's region (posted NVLink stores, no round trip), then a one-warp kernel publishes "epoch e is
// complete" into every rank's flag [its rank].  A receiver works on LOCAL memory only: it waits until its
// N flags have reached the epoch, then adds its N slots in rank order - the same order on every rank, so
// the replicated dual entries of the shared constraints stay bit-identical across ranks.  No collective
// kernel is launched, nothing waits for another GPU's kernel boundary, and the private rows' row-stream
// launch sits between the push and the wait, so the transfer hides behind it.
//
// Slots are double-buffered by epoch parity.  Signal and wait are unconditional (a trial step that the
// device-side loop control has switched off still signals and waits), so two ranks are never more than
// one epoch apart and a writer of epoch e + 1 cannot touch the half a slow reader of epoch e is using.
#ifndef PDLP_B200_PEER_CUH_
#define PDLP_B200_PEER_CUH_

#include "pdlp_peer_view.h"

namespace pdlp_b200 {

#ifdef PDLP_EMULATE
// Host stand-ins (tests only): ranks may be host threads of one process (tests/emu_lib), so the flags keep their
// release / acquire meaning.
inline void PeerFenceSystem() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
inline unsigned long long PeerLoadFlag(const unsigned long long* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
inline void PeerStoreFlag(unsigned long long* p, unsigned long long v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
inline double PeerLoadSlot(const double* p) { return *reinterpret_cast<const volatile double*>(p); }
#else
__device__ __forceinline__ void PeerFenceSystem() { __threadfence_system(); }
__device__ __forceinline__ unsigned long long PeerLoadFlag(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void PeerStoreFlag(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Slots are written by other GPUs: read them from L2, never from a stale L1 line.
__device__ __forceinline__ double PeerLoadSlot(const double* p) { return __ldcg(p); }
#endif
#ifdef PDLP_EMULATE
inline unsigned int PeerCountUp(unsigned int* p) { return (*p)++; }  // emulated CTAs run one after the other
#else
__device__ __forceinline__ unsigned int PeerCountUp(unsigned int* p) { return atomicAdd(p, 1u); }
#endif

// Blocks the calling thread until every rank's flag in MY region has reached `epoch`.
__device__ __forceinline__ void PeerWait(const unsigned long long* my_flags, int world, unsigned long long epoch) {
  for (int r = 0; r < world; ++r) {
    while (PeerLoadFlag(my_flags + r) < epoch) {
#if !defined(PDLP_EMULATE)
      __nanosleep(64);
#endif
    }
  }
}

// Publishes "everything this rank stored for row epoch `epoch` is complete" to every rank.  Launched
// behind the pushing kernel on the same stream (its stores are complete at the kernel boundary; the fence
// orders them before the flag for observers on other GPUs).
static __global__ void peer_signal_rows_kernel(PeerView view, unsigned long long epoch) {
  PeerFenceSystem();
  if (threadIdx.x < static_cast<unsigned>(view.world)) PeerStoreFlag(view.RowFlags(threadIdx.x) + view.rank, epoch);
}

// Scalar exchange, sender: `count` (<= kPeerSumsPerRank) doubles of this rank to every rank, then the flag.
// One thread; called by the kernel that produced the values.
__device__ __forceinline__ void PeerPushSums(const PeerView& view, unsigned long long epoch, const double* values, int count) {
  const int parity = static_cast<int>(epoch & 1);
  for (int r = 0; r < view.world; ++r) {
    double* slot = view.SumSlot(r, parity, view.rank);
    for (int i = 0; i < count; ++i) slot[i] = values[i];
  }
  PeerFenceSystem();
  for (int r = 0; r < view.world; ++r) PeerStoreFlag(view.SumFlags(r) + view.rank, epoch);
}
// Scalar exchange, receiver: waits, then out[i] = sum (or max) over ranks, in rank order.  One thread.
__device__ __forceinline__ void PeerCombineSums(const PeerView& view, unsigned long long epoch, double* out, int count, bool max) {
  PeerWait(view.SumFlags(view.rank), view.world, epoch);
  const int parity = static_cast<int>(epoch & 1);
  for (int i = 0; i < count; ++i) {
    double a = PeerLoadSlot(view.SumSlot(view.rank, parity, 0) + i);
    for (int r = 1; r < view.world; ++r) {
      const double b = PeerLoadSlot(view.SumSlot(view.rank, parity, r) + i);
      a = max ? fmax(a, b) : a + b;
    }
    out[i] = a;
  }
}
// The same exchange by a whole CTA (every thread calls; `values` and `count` identical in all threads): thread r serves
// destination / source r, so the eight release stores - each a system-scope fence - and the eight acquire polls run
// side by side instead of one after the other on a single thread.
__device__ __forceinline__ void PeerPushSumsCta(const PeerView& view, unsigned long long epoch, const double* values, int count) {
  const int r = threadIdx.x;
  if (r < view.world) {
    double* slot = view.SumSlot(r, static_cast<int>(epoch & 1), view.rank);
    for (int i = 0; i < count; ++i) slot[i] = values[i];
    PeerFenceSystem();
    PeerStoreFlag(view.SumFlags(r) + view.rank, epoch);
  }
}
// out: shared memory, `count` doubles, valid for every thread on return.
__device__ __forceinline__ void PeerCombineSumsCta(const PeerView& view, unsigned long long epoch, double* out, int count, bool max) {
  const int t = threadIdx.x;
  if (t < view.world) {
    while (PeerLoadFlag(view.SumFlags(view.rank) + t) < epoch) {
#if !defined(PDLP_EMULATE)
      __nanosleep(64);
#endif
    }
  }
  __syncthreads();
  if (t < count) {
    const int parity = static_cast<int>(epoch & 1);
    double v[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r) {
      if (r < view.world) v[r] = PeerLoadSlot(view.SumSlot(view.rank, parity, r) + t);
    }
    double a = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r) {
      if (r < view.world) a = max ? fmax(a, v[r]) : a + v[r];
    }
    out[t] = a;
  }
  __syncthreads();
}
static __global__ void peer_combine_sums_kernel(PeerView view, unsigned long long epoch, double* out, int count, int max) {
  if (threadIdx.x == 0) PeerCombineSums(view, epoch, out, count, max != 0);
}

// In-kernel signal of a split launch (RowBlocks::split_ctas pushing CTAs): called by every thread of a pushing CTA when
// it has stored its last row (also when the op is switched off: signals are unconditional).  Each thread fences its own
// stores system-wide, the CTA counts itself done, and the LAST one to do so publishes "row epoch complete" to every
// rank - while the other CTAs of the same launch are still streaming the private rows, so the flags are on their way
// long before the launch ends and the finishing kernel finds them set.
__device__ __forceinline__ void PeerSharedCtaDone(const PeerView& view, unsigned long long epoch, int pushing_ctas) {
  __shared__ int last_pushing_cta;
  PeerFenceSystem();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = view.DoneCounter();
    last_pushing_cta = PeerCountUp(counter) + 1u == static_cast<unsigned int>(pushing_ctas) ? 1 : 0;
    if (last_pushing_cta) *counter = 0u;  // for the next launch on this stream
  }
  __syncthreads();
  if (last_pushing_cta != 0 && static_cast<int>(threadIdx.x) < view.world) {  // one thread per destination: the release stores run side by side
    PeerFenceSystem();
    PeerStoreFlag(view.RowFlags(threadIdx.x) + view.rank, epoch);
  }
}

// The sums of rows [r0, r0 + nrows) of one work block into slot [rank] of every rank's region, by the whole CTA:
// consecutive threads store consecutive rows, one destination after the other.
__device__ __forceinline__ void PeerBlockPush(const PeerView& view, int parity, int r0, int nrows, const double* sums) {
#pragma unroll 1
  for (int r = 0; r < view.world; ++r) {
    double* dst = view.RowSlot(r, parity, view.rank) + r0;
    for (int i = threadIdx.x; i < nrows; i += kBlock) dst[i] = sums[i];
  }
}

// Phase 1 of a sharded CSR pass in peer mode: this rank's part of every shared row sum, pushed into slot
// [rank] of every rank's region.  Same interface as PartialRowOp (pdlp_skeletons.cuh).
template <class Op>
struct PeerPartialRowOp {
  static constexpr int kNumSlots = 0;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  Op op;
  PeerView view;
  int parity;
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
  __device__ void RowFinish(int row, double acc, Slots<0>&) const {  // a row longer than a work block: one thread
#pragma unroll 1
    for (int r = 0; r < view.world; ++r) view.RowSlot(r, parity, view.rank)[row] = acc;
  }
  __device__ bool PushesRow(int) const { return true; }
  __device__ void BlockPush(int r0, int nrows, const double* sums) const { PeerBlockPush(view, parity, r0, nrows, sums); }
  __device__ void PushingCtaDone(int) const {}  // never launched split: a signal kernel follows
};

// Phase 1 and the private rows in ONE row-stream launch (a CSR block cut into shared rows [0, cut_row) and
// private rows): shared rows push their partial sums like PeerPartialRowOp, private rows run the op's own
// epilogue.  The persistent grid takes work blocks in index order, shared rows first, so their NVLink stores
// drain while the same CTAs work through the private rows; as two launches the pushing kernel had to wait
// for its last store to be acknowledged before the private rows could start.
template <class Op, bool kHasPre = HasRowPrefetch<Op>::value>
struct PeerCutRowOp;
template <class Op>
struct PeerCutRowOpBase {
  static constexpr int kNumSlots = Op::kNumSlots;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return Op::SlotKindOf(s); }
  Op op;
  PeerView view;
  int parity;
  int cut_row;
  unsigned long long epoch;
  __device__ bool Begin() { return op.Begin(); }
  __device__ void PushingCtaDone(int pushing_ctas) const { PeerSharedCtaDone(view, epoch, pushing_ctas); }
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
  __device__ void Push(int row, double acc) const {  // a row longer than a work block: one thread
#pragma unroll 1
    for (int r = 0; r < view.world; ++r) view.RowSlot(r, parity, view.rank)[row] = acc;
  }
  __device__ bool PushesRow(int row) const { return row < cut_row; }
  __device__ void BlockPush(int r0, int nrows, const double* sums) const { PeerBlockPush(view, parity, r0, nrows, sums); }
};
template <class Op>
struct PeerCutRowOp<Op, true> : PeerCutRowOpBase<Op> {
  using Pre = typename Op::Pre;
  __device__ Pre RowPrefetch(int row) const { return row >= this->cut_row ? this->op.RowPrefetch(row) : Pre{}; }
  __device__ void RowFinish(int row, double acc, const Pre& pre, Slots<Op::kNumSlots>& red) const {
    if (row < this->cut_row) this->Push(row, acc);
    else this->op.RowFinish(row, acc, pre, red);
  }
};
template <class Op>
struct PeerCutRowOp<Op, false> : PeerCutRowOpBase<Op> {
  __device__ void RowFinish(int row, double acc, Slots<Op::kNumSlots>& red) const {
    if (row < this->cut_row) this->Push(row, acc);
    else this->op.RowFinish(row, acc, red);
  }
};

// Phase 2 in peer mode: wait for every rank's push, add the N local slots in rank order (or take their
// maximum), run the op's epilogue.  A map op over the shared rows.  The wait runs even when the op is
// switched off (Begin() false), so that ranks stay within one epoch of each other.
template <class Op>
struct PeerFinishRowOp {
  static constexpr int kNumSlots = Op::kNumSlots;
  __host__ __device__ static constexpr int SlotKindOf(int s) { return Op::SlotKindOf(s); }
  Op op;
  PeerView view;
  unsigned long long epoch;
  int combine_max;
  const double* slots;  // resolved by Begin(): my region, this epoch's half, source 0
  __device__ bool Begin() {
    // One thread per CTA polls the flags (system-scope acquire loads are slow, and 300 000 threads polling the same
    // eight words is what the finishing kernel spent its time on); the barrier hands the ordering to the others.
    // Every thread of the CTA calls Begin() at the top of the kernel, so the barrier is uniform.
    if (threadIdx.x == 0) PeerWait(view.RowFlags(view.rank), view.world, epoch);
    __syncthreads();
    slots = view.RowSlot(view.rank, static_cast<int>(epoch & 1), 0);
    return op.Begin();
  }
  __device__ void Apply(int64_t i, Slots<Op::kNumSlots>& acc) const {
    // All ranks' sums of the row in flight together, then added in rank order.
    double v[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r) {
      if (r < view.world) v[r] = PeerLoadSlot(slots + static_cast<size_t>(r) * static_cast<size_t>(view.cap_rows) + i);
    }
    double a = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r) {
      if (r < view.world) a = combine_max ? fmax(a, v[r]) : a + v[r];
    }
    const int row = static_cast<int>(i);
    OpRowFinish(op, row, a, OpRowPrefetch(op, row), acc);
  }
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_PEER_CUH_
