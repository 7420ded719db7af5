// Internal launch helpers shared by the op translation units.
#ifndef PDLP_B200_LAUNCH_CUH_
#define PDLP_B200_LAUNCH_CUH_

#include "pdlp_ops.cuh"

namespace pdlp_b200 {

// Base for ops that are always active and reduce nothing.
struct AlwaysActive {
  __device__ bool Begin() { return true; }
};
struct NoSlots {
  static constexpr int kNumSlots = 0;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
};

// Enqueue only (no finalize, no sync): partial sums land in ws.partials[region].
template <class Op>
inline int EnqueueMap(Workspace& ws, const Op& op, int64_t count, int region = 0) {
  const int grid = GridFor(count);
  map_reduce_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, count, ws.partials[region].get());
  ++ws.launches;
  return grid;
}
template <class Op>
inline int EnqueueRows(Workspace& ws, const Op& op, const SparseRows& s, int region = 0) {
  const int grid = s.Grid();
  rowstream_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, s.View(), ws.partials[region].get());
  ++ws.launches;
  return grid;
}

// Finalize `region` into ws.totals[offset ...] (device), no sync.
template <class Op>
inline void EnqueueFinalize(Workspace& ws, int grid, int region, int offset) {
  if (Op::kNumSlots == 0) return;
  finalize_kernel<Op><<<1, kBlock, 0, ws.stream>>>(ws.partials[region].get(), grid, ws.totals.get() + offset);
  ++ws.launches;
}

// Copies `count` totals to the pinned host buffer and waits.
inline const double* FetchTotals(Workspace& ws, int count) {
  PDLP_CUDA_CHECK(cudaMemcpyAsync(ws.host_totals, ws.totals.get(), count * sizeof(double), cudaMemcpyDeviceToHost, ws.stream));
  PDLP_CUDA_CHECK(cudaGetLastError());
  ws.Sync();
  return ws.host_totals;
}

// Run + reduce + read back.
template <class Op>
inline const double* RunMap(Workspace& ws, const Op& op, int64_t count) {
  const int grid = EnqueueMap(ws, op, count, 0);
  EnqueueFinalize<Op>(ws, grid, 0, 0);
  if (Op::kNumSlots == 0) return nullptr;
  return FetchTotals(ws, Op::kNumSlots);
}
template <class Op>
inline const double* RunRows(Workspace& ws, const Op& op, const SparseRows& s) {
  const int grid = EnqueueRows(ws, op, s, 0);
  EnqueueFinalize<Op>(ws, grid, 0, 0);
  if (Op::kNumSlots == 0) return nullptr;
  return FetchTotals(ws, Op::kNumSlots);
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_LAUNCH_CUH_
