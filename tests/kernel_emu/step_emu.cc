// CPU run of the adaptive-rule trial step (K1 - K4 and the averaging flush of
// or-tools_b200/csrc/pdlp_step_kernels.cuh) under the shim in cuda_emu.h, for
// tests/test_kernel_emulation.py.  The launch sequence and the initial StepState restate
// EnqueueAdaptiveTrials (pdlp_step.cu) and pdlp_b200_trial_steps (pdlp_capi.cu) for one GPU.
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_step_kernels.cuh"

#include <limits>

namespace {
using namespace pdlp_b200;

struct HostRows {
  const int* ptr; const int* idx; const double* val; const int* block_start; int num_blocks; int rows;
  std::vector<int> block_nz;
  RowBlocks View() {
    block_nz.resize(num_blocks + 1);
    for (int b = 0; b <= num_blocks; ++b) block_nz[b] = ptr[block_start[b]];
    return RowBlocks{ptr, block_start, block_nz.data(), num_blocks, rows};
  }
};

template <class Op>
int LaunchMap(const Op& op, int64_t count, double* partials) {
  const int grid = GridFor(count);
  cuda_emu::Launch(grid, kBlock, [=]() { map_reduce_kernel<Op>(op, count, partials); });
  return grid;
}
template <class Op>
int LaunchRows(const Op& op, const RowBlocks& rb, int max_grid, double* partials) {
  const int grid = std::max(1, std::min(rb.num_blocks, max_grid));
  cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<Op>(op, rb, partials); });
  return grid;
}
}  // namespace

// The interrupt flag of the step state (pdlp_step_state.h): raised before trial `g_interrupt_at` of the next batch
// (-1: the batch has no flag), so that the decision kernel's reaction can be followed trial by trial.
static int g_interrupt_at = -1;
static int g_interrupt_flag = 0;
extern "C" void emu_set_interrupt_at(int trial) { g_interrupt_at = trial; }

// x, y: in = start point, out = current iterate after the last accepted step.  q may be null (LP).
// state_out = {step_size, iterations_completed, num_rejected_steps, last_movement, last_nonlinearity,
//              avg_weight_sum, active, total_trials}.
extern "C" int emu_adaptive_trials(int n, int m, const int* csr_ptr, const int* csr_idx, const double* csr_val, const int* csr_blocks, int csr_num_blocks,
                                   const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                                   const double* q, const double* vlb, const double* vub, const double* clb, const double* cub, double* x, double* y,
                                   double* avg_x, double* avg_y, double step_size, double primal_weight, int trials, int constant_step, int max_grid,
                                   int defer_async, double* state_out) {
  cuda_emu::defer_async = defer_async != 0;
  HostRows csr{csr_ptr, csr_idx, csr_val, csr_blocks, csr_num_blocks, m, {}};
  HostRows csc{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  const RowBlocks by_row = csr.View(), by_col = csc.View();
  std::vector<double> xb[2] = {std::vector<double>(x, x + n), std::vector<double>(n, 0.0)};
  std::vector<double> yb[2] = {std::vector<double>(y, y + m), std::vector<double>(m, 0.0)};
  std::vector<double> kty[2] = {std::vector<double>(n, 0.0), std::vector<double>(n, 0.0)};
  std::vector<double> xbar(n, 0.0), ax(n, 0.0), ay(m, 0.0);
  for (int j = 0; j < n; ++j) {  // K^T y of the start point (pdlp_capi.cu: RowDot over the CSC copy)
    double a = 0.0;
    for (int k = csc_ptr[j]; k < csc_ptr[j + 1]; ++k) a += csc_val[k] * y[csc_idx[k]];
    kty[0][j] = a;
  }
  StepState state{};
  state.active = 1;
  state.target_iteration = 2147483647;
  state.step_size = step_size;
  state.primal_weight = primal_weight;
  state.ratio_last_two_step_sizes = 1.0;
  state.kkt_pass_limit = std::numeric_limits<double>::infinity();
  state.passes_per_rejected_step = 1.0;
  state.reduction_exponent = 0.3;
  state.growth_exponent = 0.6;
  state.mp_tau = -1.0;
  state.mp_new_tau = -1.0;
  StepState* st = &state;
  g_interrupt_flag = 0;
  if (g_interrupt_at >= 0) state.interrupt_flag = &g_interrupt_flag;

  std::vector<double> part[3] = {std::vector<double>(kMaxGrid * 2), std::vector<double>(kMaxGrid * 2), std::vector<double>(kMaxGrid * 2)};
  double* p0 = part[0].data(); double* p1 = part[1].data(); double* p2 = part[2].data();
  PrimalStepOp k1{{}, st, xb[0].data(), xb[1].data(), kty[0].data(), kty[1].data(), c, q, vlb, vub, xbar.data(), 1, ax.data()};
  DualStepOp k2{{}, st, csr_val, csr_idx, xbar.data(), yb[0].data(), yb[1].data(), clb, cub, 0, ay.data()};
  KtyStepOp k3{st, csc_val, csc_idx, yb[0].data(), yb[1].data(), xb[0].data(), xb[1].data(), kty[0].data(), kty[1].data()};
  AverageOp flush{{}, st, xb[0].data(), xb[1].data(), yb[0].data(), yb[1].data(), ax.data(), ay.data(), n, 1};
  for (int t = 0; t < trials; ++t) {
    if (t == g_interrupt_at) g_interrupt_flag = 1;
    const int g1 = LaunchMap(k1, n, p0);
    const int g2 = LaunchRows(k2, by_row, max_grid, p1);
    const int g3 = LaunchRows(k3, by_col, max_grid, p2);
    cuda_emu::Launch(1, kBlock, [=]() { decide_adaptive_kernel(st, p0, g1, p1, g2, p2, g3, constant_step ? 1 : 0, PeerView{}, 0ULL); });
  }
  if (trials > 0) {
    LaunchMap(flush, static_cast<int64_t>(n) + m, p0);
    cuda_emu::Launch(1, 1, [=]() { finish_average_kernel(st); });
  }
  std::copy(xb[state.cur].begin(), xb[state.cur].end(), x);
  std::copy(yb[state.cur].begin(), yb[state.cur].end(), y);
  std::copy(ax.begin(), ax.end(), avg_x);
  std::copy(ay.begin(), ay.end(), avg_y);
  state_out[0] = state.step_size; state_out[1] = state.iterations_completed; state_out[2] = state.num_rejected_steps;
  state_out[3] = state.last_movement; state_out[4] = state.last_nonlinearity; state_out[5] = state.avg_weight_sum;
  state_out[6] = state.active; state_out[7] = static_cast<double>(state.total_trials);
  state_out[8] = state.stopped_by_interrupt;
  return 0;
}

// Malitsky-Pock rule: the launch sequence of EnqueueMalitskyPockTrials (pdlp_step.cu) for one GPU.
// params = {step_size_downscaling_factor, linesearch_contraction_factor, step_size_interpolation}.
// state_out as above, with [5] = primal average weight sum, [8] = dual average weight sum,
// [9] = ratio_last_two_step_sizes.
extern "C" int emu_malitsky_pock_trials(int n, int m, const int* csr_ptr, const int* csr_idx, const double* csr_val, const int* csr_blocks, int csr_num_blocks,
                                        const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                                        const double* q, const double* vlb, const double* vub, const double* clb, const double* cub, double* x, double* y,
                                        double* avg_x, double* avg_y, double step_size, double primal_weight, const double* params, int trials, int max_grid,
                                        double* state_out) {
  cuda_emu::defer_async = false;
  HostRows csr{csr_ptr, csr_idx, csr_val, csr_blocks, csr_num_blocks, m, {}};
  HostRows csc{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  const RowBlocks by_row = csr.View(), by_col = csc.View();
  std::vector<double> xb[2] = {std::vector<double>(x, x + n), std::vector<double>(n, 0.0)};
  std::vector<double> yb[2] = {std::vector<double>(y, y + m), std::vector<double>(m, 0.0)};
  std::vector<double> kty[2] = {std::vector<double>(n, 0.0), std::vector<double>(n, 0.0)};
  std::vector<double> xbar(n, 0.0), ax(n, 0.0), ay(m, 0.0);
  for (int j = 0; j < n; ++j) {
    double a = 0.0;
    for (int k = csc_ptr[j]; k < csc_ptr[j + 1]; ++k) a += csc_val[k] * y[csc_idx[k]];
    kty[0][j] = a;
  }
  StepState state{};
  state.active = 1;
  state.target_iteration = 2147483647;
  state.step_size = step_size;
  state.primal_weight = primal_weight;
  state.ratio_last_two_step_sizes = 1.0;
  state.kkt_pass_limit = std::numeric_limits<double>::infinity();
  state.passes_per_rejected_step = 1.0;
  state.reduction_exponent = 0.3;
  state.growth_exponent = 0.6;
  state.mp_downscaling = params[0];
  state.mp_contraction = params[1];
  state.mp_interpolation = params[2];
  state.mp_tau = -1.0;
  state.mp_new_tau = -1.0;
  StepState* st = &state;
  std::vector<double> part[3] = {std::vector<double>(kMaxGrid * 2), std::vector<double>(kMaxGrid * 2), std::vector<double>(kMaxGrid * 2)};
  double* p0 = part[0].data(); double* p1 = part[1].data(); double* p2 = part[2].data();
  PrimalStepOp k1{{}, st, xb[0].data(), xb[1].data(), kty[0].data(), kty[1].data(), c, q, vlb, vub, xbar.data(), 0, nullptr};
  ExtrapolateOp kx{{}, st, xb[0].data(), xb[1].data(), xbar.data()};
  DualStepOp k2{{}, st, csr_val, csr_idx, xbar.data(), yb[0].data(), yb[1].data(), clb, cub, 0, nullptr};
  KtyStepOp k3{st, csc_val, csc_idx, yb[0].data(), yb[1].data(), xb[0].data(), xb[1].data(), kty[0].data(), kty[1].data()};
  AverageOp k5{{}, st, xb[0].data(), xb[1].data(), yb[0].data(), yb[1].data(), ax.data(), ay.data(), n, 0};
  for (int t = 0; t < trials; ++t) {
    cuda_emu::Launch(1, 1, [=]() { mp_begin_kernel(st); });
    const int g1 = LaunchMap(k1, n, p0);
    LaunchMap(kx, n, p1);
    const int g2 = LaunchRows(k2, by_row, max_grid, p1);
    const int g3 = LaunchRows(k3, by_col, max_grid, p2);
    cuda_emu::Launch(1, kBlock, [=]() { decide_mp_kernel(st, p0, g1, p1, g2, p2, g3, PeerView{}, 0ULL); });
    LaunchMap(k5, static_cast<int64_t>(n) + m, p0);
  }
  std::copy(xb[state.cur].begin(), xb[state.cur].end(), x);
  std::copy(yb[state.cur].begin(), yb[state.cur].end(), y);
  std::copy(ax.begin(), ax.end(), avg_x);
  std::copy(ay.begin(), ay.end(), avg_y);
  state_out[0] = state.step_size; state_out[1] = state.iterations_completed; state_out[2] = state.num_rejected_steps;
  state_out[3] = state.last_movement; state_out[4] = state.last_nonlinearity; state_out[5] = state.avg_weight_sum;
  state_out[6] = state.active; state_out[7] = static_cast<double>(state.total_trials);
  state_out[8] = state.avg_weight_sum_dual; state_out[9] = state.ratio_last_two_step_sizes;
  return 0;
}

// ---- column-sharded dual step (K2), one rank at a time ----
// The two device phases of pdlp_launch.cuh:EnqueueRows for a CSR block cut into shared rows
// [0, cut_row) and private rows; the test sums the ranks' partial row sums in between (the
// all-reduce).  block_start comes from pdlp_b200.work_blocks(ptr, cut_row); cut_block = number of
// work blocks before the cut (0: no shared rows, num_blocks: all shared).
namespace {
StepState FreshState(double step_size, double primal_weight) {
  StepState state{};
  state.active = 1;
  state.target_iteration = 2147483647;
  state.step_size = step_size;
  state.primal_weight = primal_weight;
  state.ratio_last_two_step_sizes = 1.0;
  state.mp_tau = -1.0;
  state.mp_new_tau = -1.0;
  return state;
}
RowBlocks ViewOfBlocks(const RowBlocks& all, int first, int count) {  // pdlp_matrix.cuh: SparseRows::ViewOfBlocks
  return RowBlocks{all.row_ptr, all.block_start + first, all.block_nz + first, count, all.rows};
}
}  // namespace

// Phase 1: this rank's part of the shared rows' sums of K xbar -> sums[cut_row].
extern "C" int emu_sharded_dual_partial(int m_local, const int* ptr, const int* idx, const double* val, const int* block_start, int num_blocks, int cut_block,
                                        const double* xbar, double* y0, double* y1, const double* clb, const double* cub, double step_size, double primal_weight,
                                        int max_grid, double* sums) {
  cuda_emu::defer_async = false;
  HostRows csr{ptr, idx, val, block_start, num_blocks, m_local, {}};
  const RowBlocks all = csr.View();
  StepState state = FreshState(step_size, primal_weight);
  if (cut_block > 0) {
    const DualStepOp op{{}, &state, val, idx, xbar, y0, y1, clb, cub, 0, nullptr};
    std::vector<double> partials(kMaxGrid, 0.0);
    LaunchRows(PartialRowOp<DualStepOp>{op, sums}, ViewOfBlocks(all, 0, cut_block), max_grid, partials.data());
  }
  return 0;
}

// Phase 2: private rows through the op's own epilogue inside the row-stream kernel, shared rows
// through FinishRowOp on the all-reduced sums.  y1 receives y'; *dy2 = this rank's part of |dy|^2
// (rows below skip_below are counted on rank 0 only).
extern "C" int emu_sharded_dual_finish(int m_local, const int* ptr, const int* idx, const double* val, const int* block_start, int num_blocks, int cut_block,
                                       int cut_row, const double* xbar, double* y0, double* y1, const double* clb, const double* cub, double step_size,
                                       double primal_weight, int skip_below, int max_grid, const double* sums, double* dy2) {
  cuda_emu::defer_async = false;
  HostRows csr{ptr, idx, val, block_start, num_blocks, m_local, {}};
  const RowBlocks all = csr.View();
  StepState state = FreshState(step_size, primal_weight);
  const DualStepOp op{{}, &state, val, idx, xbar, y0, y1, clb, cub, skip_below, nullptr};
  std::vector<double> partials(2 * kMaxGrid, 0.0);
  int grid = 0;
  if (num_blocks - cut_block > 0) grid = LaunchRows(op, ViewOfBlocks(all, cut_block, num_blocks - cut_block), max_grid, partials.data());
  if (cut_row > 0) grid += LaunchMap(FinishRowOp<DualStepOp>{op, sums}, cut_row, partials.data() + grid);
  double total = 0.0;
  for (int b = 0; b < grid; ++b) total += partials[b];
  *dy2 = total;
  return 0;
}

// ---- the same two phases through peer memory (pdlp_peer.cuh): all emulated ranks live in this process,
// `regions[r]` is rank r's region (PeerView::Bytes(world, cap_rows) bytes, zeroed by the caller once). ----
namespace {
PeerView ViewOf(int rank, int world, long long cap_rows, void* const* regions) {
  PeerView v;
  v.rank = rank;
  v.world = world;
  v.cap_rows = cap_rows;
  for (int r = 0; r < world; ++r) v.base[r] = static_cast<char*>(regions[r]);
  return v;
}
}  // namespace

extern "C" long long emu_peer_region_bytes(int world, long long cap_rows) { return static_cast<long long>(PeerView::Bytes(world, cap_rows)); }

// Phase 1 + signal of `rank` for row epoch `epoch`: pushes into every rank's region.
extern "C" int emu_peer_dual_partial(int m_local, const int* ptr, const int* idx, const double* val, const int* block_start, int num_blocks, int cut_block,
                                     const double* xbar, double* y0, double* y1, const double* clb, const double* cub, double step_size, double primal_weight,
                                     int max_grid, int rank, int world, long long cap_rows, void* const* regions, unsigned long long epoch) {
  cuda_emu::defer_async = false;
  HostRows csr{ptr, idx, val, block_start, num_blocks, m_local, {}};
  const RowBlocks all = csr.View();
  StepState state = FreshState(step_size, primal_weight);
  const PeerView view = ViewOf(rank, world, cap_rows, regions);
  if (cut_block > 0) {
    const DualStepOp op{{}, &state, val, idx, xbar, y0, y1, clb, cub, 0, nullptr};
    std::vector<double> partials(kMaxGrid, 0.0);
    LaunchRows(PeerPartialRowOp<DualStepOp>{op, view, static_cast<int>(epoch & 1)}, ViewOfBlocks(all, 0, cut_block), max_grid, partials.data());
  }
  cuda_emu::Launch(1, 32, [=]() { peer_signal_rows_kernel(view, epoch); });
  return 0;
}

// Phase 2 of `rank`: private rows in the row-stream kernel, shared rows by PeerFinishRowOp (wait + rank-order sum).
extern "C" int emu_peer_dual_finish(int m_local, const int* ptr, const int* idx, const double* val, const int* block_start, int num_blocks, int cut_block,
                                    int cut_row, const double* xbar, double* y0, double* y1, const double* clb, const double* cub, double step_size,
                                    double primal_weight, int skip_below, int max_grid, int rank, int world, long long cap_rows, void* const* regions,
                                    unsigned long long epoch, double* dy2) {
  cuda_emu::defer_async = false;
  HostRows csr{ptr, idx, val, block_start, num_blocks, m_local, {}};
  const RowBlocks all = csr.View();
  StepState state = FreshState(step_size, primal_weight);
  const DualStepOp op{{}, &state, val, idx, xbar, y0, y1, clb, cub, skip_below, nullptr};
  const PeerView view = ViewOf(rank, world, cap_rows, regions);
  std::vector<double> partials(2 * kMaxGrid, 0.0);
  int grid = 0;
  if (num_blocks - cut_block > 0) grid = LaunchRows(op, ViewOfBlocks(all, cut_block, num_blocks - cut_block), max_grid, partials.data());
  if (cut_row > 0) grid += LaunchMap(PeerFinishRowOp<DualStepOp>{op, view, epoch, 0, nullptr}, cut_row, partials.data() + grid);
  double total = 0.0;
  for (int b = 0; b < grid; ++b) total += partials[b];
  *dy2 = total;
  return 0;
}

// The scalar exchange: every rank pushes `count` values for sum epoch `epoch`, then every rank combines.
// values: world x count in, out: world x count (each rank's combined copy; must be identical rows).
extern "C" int emu_peer_sums(int world, long long cap_rows, void* const* regions, unsigned long long epoch, int count, int max, const double* values, double* out) {
  for (int r = 0; r < world; ++r) {
    const PeerView view = ViewOf(r, world, cap_rows, regions);
    const double* mine = values + static_cast<size_t>(r) * count;
    // Even epochs through the one-thread forms, odd epochs (the sums of the test) through the CTA-wide forms the step kernels use.
    if (!(epoch & 1)) cuda_emu::Launch(1, 1, [=]() { PeerPushSums(view, epoch, mine, count); });
    else cuda_emu::Launch(1, kBlock, [=]() { PeerPushSumsCta(view, epoch, mine, count); });
  }
  for (int r = 0; r < world; ++r) {
    const PeerView view = ViewOf(r, world, cap_rows, regions);
    double* dst = out + static_cast<size_t>(r) * count;
    if (!(epoch & 1)) {
      cuda_emu::Launch(1, 32, [=]() { peer_combine_sums_kernel(view, epoch, dst, count, max); });
    } else {
      cuda_emu::Launch(1, kBlock, [=]() {
        __shared__ double combined[kPeerSumsPerRank];
        PeerCombineSumsCta(view, epoch, combined, count, max != 0);
        if (static_cast<int>(threadIdx.x) < count) dst[threadIdx.x] = combined[threadIdx.x];
      });
    }
  }
  return 0;
}

// Phase 1 + private rows of `rank` in one launch (PeerCutRowOp), then the signal; `*dy2_private` = the private
// rows' part of |dy|^2.  emu_peer_dual_finish with cut_block == num_blocks then only runs the shared rows' finish.
extern "C" int emu_peer_dual_fused(int m_local, const int* ptr, const int* idx, const double* val, const int* block_start, int num_blocks, int cut_row,
                                   const double* xbar, double* y0, double* y1, const double* clb, const double* cub, double step_size, double primal_weight,
                                   int skip_below, int max_grid, int rank, int world, long long cap_rows, void* const* regions, unsigned long long epoch,
                                   double* dy2_private) {
  cuda_emu::defer_async = false;
  HostRows csr{ptr, idx, val, block_start, num_blocks, m_local, {}};
  RowBlocks all = csr.View();
  // The launcher's static split (pdlp_launch.cuh:PeerSharedCtas): the first CTAs take the shared rows' work blocks, the
  // others the private rows', concurrently.  Odd ranks run unsplit, so both block orders are exercised.
  int cut_block = 0;
  while (cut_block < num_blocks && block_start[cut_block] < cut_row) ++cut_block;
  const int grid_size = std::max(1, std::min(num_blocks, max_grid));
  if ((rank & 1) == 0 && grid_size >= 2 && cut_block > 0 && cut_block < num_blocks) {
    all.split_block = cut_block;
    all.split_ctas = std::max(1, std::min(grid_size / 2, cut_block));
  }
  StepState state = FreshState(step_size, primal_weight);
  const PeerView view = ViewOf(rank, world, cap_rows, regions);
  PeerCutRowOp<DualStepOp> op;
  op.op = DualStepOp{{}, &state, val, idx, xbar, y0, y1, clb, cub, skip_below, nullptr};
  op.view = view; op.parity = static_cast<int>(epoch & 1); op.cut_row = cut_row; op.epoch = epoch;
  std::vector<double> partials(kMaxGrid, 0.0);
  const int grid = LaunchRows(op, all, max_grid, partials.data());
  if (all.split_ctas == 0) cuda_emu::Launch(1, 32, [=]() { peer_signal_rows_kernel(view, epoch); });  // a split launch has signalled from inside
  double total = 0.0;
  for (int b = 0; b < grid; ++b) total += partials[b];
  *dy2_private = total;
  return 0;
}
