// Internal launch helpers shared by the op translation units.
//
// Column-sharded solves: every reduction names the SIDE its index range lives on
// (pdlp_ops.cuh).  Each rank accumulates only the terms it owns - all of its columns, its private
// rows, and on rank 0 the rows shared by several ranks - and every finalize is followed by an
// all-reduce of the totals.  A row-stream launch over the CSR block (`SparseRows::partial_sums`)
// runs in two phases around an all-reduce of the SHARED rows' sums: the only O(rows) collective.
#ifndef PDLP_B200_LAUNCH_CUH_
#define PDLP_B200_LAUNCH_CUH_

#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "pdlp_ops.cuh"

namespace pdlp_b200 {

inline bool IsSharded(const Workspace& ws) { return ws.comm != nullptr && ws.comm->sharded(); }

// Persistent grid of a row-stream launch: one wave of resident CTAs (SMs x occupancy of this
// instantiation, asked from the runtime once), never more CTAs than work blocks.
template <class Op>
inline int RowStreamGrid(const SparseRows& s) {
  static const int resident = [] {
    int device = 0, sms = 148, per_sm = PDLP_ROWSTREAM_MIN_CTAS;
    if (cudaGetDevice(&device) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rowstream_kernel<Op>, kBlock, 0) != cudaSuccess || per_sm < 1) per_sm = PDLP_ROWSTREAM_MIN_CTAS;
    const int g = sms * per_sm;
    return g > kMaxGrid ? kMaxGrid : g;
  }();
  return s.num_blocks < 1 ? 1 : (s.num_blocks > resident ? resident : s.num_blocks);
}

// Enqueue only (no finalize, no sync): partial sums land in ws.partials[region].
template <class Op>
inline int EnqueueMap(Workspace& ws, const Op& op, int64_t count, int region = 0) {
  const int grid = GridFor(count);
  map_reduce_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, count, ws.partials[region].get());
  ++ws.launches;
  return grid;
}
// How many CTAs of a fused peer pass take the shared rows (RowBlocks::split_ctas).  The shared part is bound by the
// NVLink stores (8 bytes per row and peer at ~60 % of 900 GB/s per direction) or by its own stream, the private part by
// HBM (~3.7 TB/s for this kernel on the whole GPU): the share f of the grid that makes both sides finish together,
// max(t_link, t_shared / f) = t_private / (1 - f).  PDLP_B200_PEER_SHARED_CTAS_PCT overrides (0 = no split).
inline int PeerSharedCtas(const SparseRows& s, int world, int grid) {
  static const int forced = [] {
    const char* env = std::getenv("PDLP_B200_PEER_SHARED_CTAS_PCT");
    return env != nullptr ? std::atoi(env) : -1;
  }();
  if (grid < 2 || s.cut_block <= 0 || s.cut_block >= s.num_blocks) return 0;
  double f;
  if (forced >= 0) {
    if (forced == 0) return 0;
    f = forced / 100.0;
  } else {
    const double shared_rows = s.cut_row, private_rows = s.rows - s.cut_row;
    const double t_link = (world - 1) * 8.0 * shared_rows / 540e9;
    const double t_shared = (12.0 * static_cast<double>(s.cut_nnz) + 16.0 * shared_rows) / 3.7e12;
    const double t_private = (12.0 * static_cast<double>(s.nnz - s.cut_nnz) + 64.0 * private_rows) / 3.7e12;
    // Link-bound shared side: t_link = t_private / (1 - f).  Stream-bound: t_shared / f = t_private / (1 - f).
    const double f_stream = t_shared / (t_shared + t_private);
    const double f_link = t_link > t_private ? 0.5 : 1.0 - t_private / std::max(t_link, 1e-12);
    f = std::max(f_stream, std::min(f_link, 0.5));
    if (!(f > 0.0)) f = f_stream;
  }
  int ctas = static_cast<int>(f * grid + 0.5);
  ctas = std::max(1, std::min(ctas, grid - 1));
  return std::min(ctas, s.cut_block);
}

// A pass over the column-sharded CSR block with the shared rows' sums exchanged through peer memory
// (pdlp_peer.cuh): push + signal, then the private rows (which hide the transfer), then wait + finish.
template <class Op>
inline int EnqueueRowsPeer(Workspace& ws, const Op& op, const SparseRows& s, int region) {
  const PeerView view = ws.comm->peer_view();
  double* partials = ws.partials[region].get();
  const bool cut = s.cut_block >= 0;
  const int shared_blocks = cut ? s.cut_block : s.num_blocks;
  const int private_blocks = cut ? s.num_blocks - s.cut_block : 0;
  const int64_t shared_rows = cut ? s.cut_row : s.rows;
  unsigned long long epoch = 0;
  int grid = 0;
  if (shared_rows > 0) epoch = ws.comm->NextRowEpoch();  // the same on every rank
  auto mark = [&]() {
    if (ws.profile_peer == nullptr) return;
    cudaEvent_t e;
    PDLP_CUDA_CHECK(cudaEventCreate(&e));
    PDLP_CUDA_CHECK(cudaEventRecord(e, ws.stream));
    ws.profile_peer->push_back(e);
  };
  mark();
  if (shared_rows > 0 && shared_blocks > 0 && private_blocks > 0) {
    // One launch over the whole block: shared rows push, private rows finish in place (PeerCutRowOp).
    using Cut = PeerCutRowOp<Op>;
    Cut cut_op;
    cut_op.op = op; cut_op.view = view; cut_op.parity = static_cast<int>(epoch & 1); cut_op.cut_row = static_cast<int>(s.cut_row); cut_op.epoch = epoch;
    grid = RowStreamGrid<Cut>(s);
    RowBlocks blocks = s.View();
    blocks.split_block = shared_blocks;
    blocks.split_ctas = PeerSharedCtas(s, view.world, grid);
    rowstream_kernel<Cut><<<grid, kBlock, 0, ws.stream>>>(cut_op, blocks, partials);
    mark();
    ++ws.launches;
    if (blocks.split_ctas == 0) {  // a split launch signals from inside, as soon as its pushing CTAs are done
      peer_signal_rows_kernel<<<1, 32, 0, ws.stream>>>(view, epoch);
      ++ws.launches;
    }
    mark();
  } else {
    if (shared_rows > 0) {
      if (shared_blocks > 0) {
        const RowBlocks v = s.ViewOfBlocks(0, shared_blocks);
        const int g = std::min(shared_blocks, RowStreamGrid<PeerPartialRowOp<Op>>(s));
        rowstream_kernel<PeerPartialRowOp<Op>><<<g, kBlock, 0, ws.stream>>>(PeerPartialRowOp<Op>{op, view, static_cast<int>(epoch & 1)}, v, partials);
        ++ws.launches;
      }
      peer_signal_rows_kernel<<<1, 32, 0, ws.stream>>>(view, epoch);
      ++ws.launches;
    }
    if (private_blocks > 0) {
      const RowBlocks v = s.ViewOfBlocks(s.cut_block, private_blocks);
      grid = std::min(private_blocks, RowStreamGrid<Op>(s));
      rowstream_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, v, partials);
      ++ws.launches;
    }
  }
  if (shared_rows > 0) {
    const int g = GridFor(shared_rows);
    map_reduce_kernel<PeerFinishRowOp<Op>><<<g, kBlock, 0, ws.stream>>>(PeerFinishRowOp<Op>{op, view, epoch, RowCombineIsMax(op) ? 1 : 0, nullptr}, shared_rows,
                                                                      partials + static_cast<size_t>(grid) * Op::kNumSlots);
    ++ws.launches;
    grid += g;
  }
  mark();
  return grid > 0 ? grid : 1;
}

template <class Op>
inline int EnqueueRows(Workspace& ws, const Op& op, const SparseRows& s, int region = 0) {
  if (s.partial_sums && IsSharded(ws)) {
    if (ws.peer) return EnqueueRowsPeer(ws, op, s, region);
    double* sums = ws.exchange_rows;
    double* partials = ws.partials[region].get();
    if (s.cut_block >= 0) {
      // The block's rows are cut into shared rows [0, cut_row) and rows private to this rank.  Only the
      // shared rows need the exchange: their partial sums are stored and all-reduced, then finished by a
      // map kernel over cut_row rows; the private rows run the op's own epilogue inside the row-stream
      // kernel, exactly as on a single GPU (no partial-sum round trip, no second pass over their vectors).
      // Reduction partials of the two finishing launches sit behind one another in the region's table.
      const int shared_blocks = s.cut_block, private_blocks = s.num_blocks - s.cut_block;
      if (shared_blocks > 0) {
        const RowBlocks v = s.ViewOfBlocks(0, shared_blocks);
        const int g = std::min(shared_blocks, RowStreamGrid<PartialRowOp<Op>>(s));
        rowstream_kernel<PartialRowOp<Op>><<<g, kBlock, 0, ws.stream>>>(PartialRowOp<Op>{op, sums}, v, partials);
        ++ws.launches;
        ws.comm->AllReduce(sums, static_cast<size_t>(s.cut_row), RowCombineIsMax(op), ws.stream);
      }
      int grid = 0;
      if (private_blocks > 0) {
        const RowBlocks v = s.ViewOfBlocks(s.cut_block, private_blocks);
        grid = std::min(private_blocks, RowStreamGrid<Op>(s));
        rowstream_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, v, partials);
        ++ws.launches;
      }
      if (s.cut_row > 0) {
        const int g = GridFor(s.cut_row);
        map_reduce_kernel<FinishRowOp<Op>><<<g, kBlock, 0, ws.stream>>>(FinishRowOp<Op>{op, sums}, s.cut_row, partials + static_cast<size_t>(grid) * Op::kNumSlots);
        ++ws.launches;
        grid += g;
      }
      return grid > 0 ? grid : 1;
    }
    rowstream_kernel<PartialRowOp<Op>><<<RowStreamGrid<PartialRowOp<Op>>(s), kBlock, 0, ws.stream>>>(PartialRowOp<Op>{op, sums}, s.View(), ws.partials[region].get());
    ws.comm->AllReduce(sums, static_cast<size_t>(ws.shared_rows), RowCombineIsMax(op), ws.stream);
    const int grid = GridFor(s.rows);
    map_reduce_kernel<FinishRowOp<Op>><<<grid, kBlock, 0, ws.stream>>>(FinishRowOp<Op>{op, sums}, s.rows, ws.partials[region].get());
    ws.launches += 2;
    return grid;
  }
  const int grid = RowStreamGrid<Op>(s);
  rowstream_kernel<Op><<<grid, kBlock, 0, ws.stream>>>(op, s.View(), ws.partials[region].get());
  ++ws.launches;
  return grid;
}
// Side of the reduction slots of a row-stream launch: the CSC block's "rows" are this rank's
// columns of K, the CSR block's rows its dual side.
inline Side SideOfRows(const SparseRows& s) { return s.sharded_rows ? Side::kSharded : Side::kDual; }

template <class Op>
__global__ void select_max_slots_kernel(double* totals, const double* max_reduced) {
  const int s = threadIdx.x;
  if (s < Op::kNumSlots && Op::SlotKindOf(s) == kMax) totals[s] = max_reduced[s];
}

// Finalize `region` into ws.totals[offset ...] (device) and, for sharded sides, combine the
// per-rank totals in place; no sync.
template <class Op>
inline void EnqueueFinalize(Workspace& ws, int grid, int region, int offset, Side side) {
  if (Op::kNumSlots == 0) return;
  double* totals = ws.totals.get() + offset;
  finalize_kernel<Op><<<1, kBlock, 0, ws.stream>>>(ws.partials[region].get(), grid, totals);
  ++ws.launches;
  (void)side;
  if (!IsSharded(ws)) return;
  bool any_max = false, any_sum = false;
  for (int s = 0; s < Op::kNumSlots; ++s) (Op::SlotKindOf(s) == kMax ? any_max : any_sum) = true;
  if (!any_max) {
    ws.comm->AllReduce(totals, Op::kNumSlots, false, ws.stream);
  } else if (!any_sum) {
    ws.comm->AllReduce(totals, Op::kNumSlots, true, ws.stream);
  } else {
    double* tmp = ws.combine_tmp.get();
    PDLP_CUDA_CHECK(cudaMemcpyAsync(tmp, totals, Op::kNumSlots * sizeof(double), cudaMemcpyDeviceToDevice, ws.stream));
    ws.comm->AllReduce(totals, Op::kNumSlots, false, ws.stream);
    ws.comm->AllReduce(tmp, Op::kNumSlots, true, ws.stream);
    select_max_slots_kernel<Op><<<1, 32, 0, ws.stream>>>(totals, tmp);
    ++ws.launches;
  }
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
inline const double* RunMap(Workspace& ws, const Op& op, int64_t count, Side side) {
  const int grid = EnqueueMap(ws, op, count, 0);
  EnqueueFinalize<Op>(ws, grid, 0, 0, side);
  if (Op::kNumSlots == 0) return nullptr;
  return FetchTotals(ws, Op::kNumSlots);
}
template <class Op>
inline const double* RunRows(Workspace& ws, const Op& op, const SparseRows& s) {
  const int grid = EnqueueRows(ws, op, s, 0);
  EnqueueFinalize<Op>(ws, grid, 0, 0, SideOfRows(s));
  if (Op::kNumSlots == 0) return nullptr;
  return FetchTotals(ws, Op::kNumSlots);
}

}  // namespace pdlp_b200
#endif  // PDLP_B200_LAUNCH_CUH_
