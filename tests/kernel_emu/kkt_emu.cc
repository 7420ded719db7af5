// CPU run of the termination-statistics ops (or-tools_b200/csrc/pdlp_kkt_kernels.cuh) under the shim
// in cuda_emu.h, for tests/test_kernel_emulation.py: the two fused passes of ComputeKktTotals
// (pdlp_kkt.cu) for one GPU, and the reduced-cost pass.
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_kkt_kernels.cuh"

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
void RunRows(const Op& op, const RowBlocks& rb, int max_grid, double* totals) {
  const int grid = std::max(1, std::min(rb.num_blocks, max_grid));
  std::vector<double> partials(static_cast<size_t>(grid) * std::max(1, Op::kNumSlots), 0.0);
  double* p = partials.data();
  cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<Op>(op, rb, p); });
  if constexpr (Op::kNumSlots > 0) cuda_emu::Launch(1, kBlock, [=]() { finalize_kernel<Op>(p, grid, totals); });
}
}  // namespace

// row_totals[9], col_totals[15]: the reduction slots of KktRowOp / KktColOp in their order.
extern "C" int emu_kkt_totals(int n, int m, const int* csr_ptr, const int* csr_idx, const double* csr_val, const int* csr_blocks, int csr_num_blocks,
                              const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                              const double* q, const double* vlb, const double* vub, const double* clb, const double* cub, const double* row_scale,
                              const double* col_scale, const double* x, const double* y, const double* x_start, const double* y_start, double cw_primal_offset,
                              double cw_dual_offset, int flag, int max_grid, double* row_totals, double* col_totals) {
  cuda_emu::defer_async = false;
  HostRows csr{csr_ptr, csr_idx, csr_val, csr_blocks, csr_num_blocks, m, {}};
  HostRows csc{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  RunRows(KktRowOp{{}, csr_val, csr_idx, x, y, y_start, row_scale, clb, cub, cw_primal_offset, 0}, csr.View(), max_grid, row_totals);
  RunRows(KktColOp{{}, csc_val, csc_idx, y, x, x_start, col_scale, c, q, vlb, vub, cw_dual_offset, flag}, csc.View(), max_grid, col_totals);
  return 0;
}

extern "C" int emu_reduced_costs(int n, const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                                 const double* q, const double* vlb, const double* vub, const double* x, const double* y, int zero_objective, int flag,
                                 int max_grid, double* rc) {
  cuda_emu::defer_async = false;
  HostRows csc{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  RunRows(ReducedCostOp{{}, {}, csc_val, csc_idx, y, x, c, q, vlb, vub, rc, zero_objective, flag}, csc.View(), max_grid, nullptr);
  return 0;
}

// The split form of the same statistics (pdlp_kkt.cu: ComputeProducts + ComputeKktTotalsFromProducts): K x and K^T y
// of THREE points - (x, y), (x_other, y_other), (x_start, y_start) - by rowstream_multi_kernel with one stream of each
// matrix copy, then the two dense epilogue passes for the first point.  kx_out: 3 m doubles, kty_out: 3 n doubles.
extern "C" int emu_kkt_totals_split(int n, int m, const int* csr_ptr, const int* csr_idx, const double* csr_val, const int* csr_blocks, int csr_num_blocks,
                                    const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                                    const double* q, const double* vlb, const double* vub, const double* clb, const double* cub, const double* row_scale,
                                    const double* col_scale, const double* x, const double* y, const double* x_start, const double* y_start, double cw_primal_offset,
                                    double cw_dual_offset, int flag, int max_grid, const double* x_other, const double* y_other, double* row_totals, double* col_totals,
                                    double* kx_out, double* kty_out) {
  cuda_emu::defer_async = false;
  HostRows csr{csr_ptr, csr_idx, csr_val, csr_blocks, csr_num_blocks, m, {}};
  HostRows csc{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  const RowBlocks by_row = csr.View(), by_col = csc.View();
  MultiDotOp<3> rows{{}, csr_val, csr_idx, {x, x_other, x_start}, {kx_out, kx_out + m, kx_out + 2 * m}};
  MultiDotOp<3> cols{{}, csc_val, csc_idx, {y, y_other, y_start}, {kty_out, kty_out + n, kty_out + 2 * n}};
  const int grid_r = std::max(1, std::min(by_row.num_blocks, max_grid)), grid_c = std::max(1, std::min(by_col.num_blocks, max_grid));
  cuda_emu::Launch(grid_r, kBlock, [=]() { rowstream_multi_kernel<MultiDotOp<3>>(rows, by_row); });
  cuda_emu::Launch(grid_c, kBlock, [=]() { rowstream_multi_kernel<MultiDotOp<3>>(cols, by_col); });
  auto run_map = [&](auto op, int64_t count, double* totals) {
    using Op = decltype(op);
    const int grid = GridFor(count);
    std::vector<double> partials(static_cast<size_t>(grid) * Op::kNumSlots, 0.0);
    double* p = partials.data();
    cuda_emu::Launch(grid, kBlock, [=]() { map_reduce_kernel<Op>(op, count, p); });
    cuda_emu::Launch(1, kBlock, [=]() { finalize_kernel<Op>(p, grid, totals); });
  };
  run_map(KktRowMapOp{{}, KktRowOp{{}, csr_val, csr_idx, x, y, y_start, row_scale, clb, cub, cw_primal_offset, 0}, kx_out}, m, row_totals);
  run_map(KktColMapOp{{}, KktColOp{{}, csc_val, csc_idx, y, x, x_start, col_scale, c, q, vlb, vub, cw_dual_offset, flag}, kty_out}, n, col_totals);
  return 0;
}
