// CPU run of the rescaling ops (or-tools_b200/csrc/pdlp_ops_kernels.cuh) under the shim in cuda_emu.h:
// the launch sequence of PreprocessSolver::ApplyRescaling (pdlp_preprocess.cu) with ScaledRowNorms,
// DivideBySqrtOfDivisor, ScaleProblemVectors and ScaleMatrixValues (pdlp_ops.cu) for one GPU.
// For tests/test_kernel_emulation.py.
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_ops_kernels.cuh"

namespace {
using namespace pdlp_b200;

struct HostRows {
  const int* ptr; const int* idx; double* val; const int* block_start; int num_blocks; int rows;
  std::vector<int> block_nz;
  RowBlocks View() {
    block_nz.resize(num_blocks + 1);
    for (int b = 0; b <= num_blocks; ++b) block_nz[b] = ptr[block_start[b]];
    return RowBlocks{ptr, block_start, block_nz.data(), num_blocks, rows};
  }
};

template <class Op>
void Rows(const Op& op, const RowBlocks& rb, int max_grid) {
  const int grid = std::max(1, std::min(rb.num_blocks, max_grid));
  cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<Op>(op, rb, nullptr); });
}
template <class Op>
void Map(const Op& op, int64_t count) {
  if (count <= 0) return;
  cuda_emu::Launch(GridFor(count), kBlock, [=]() { map_reduce_kernel<Op>(op, count, nullptr); });
}
void ScaleMatrix(HostRows& s, const double* idx_scale, const double* row_scale) {
  if (s.rows == 0) return;
  const RowBlocks rb = s.View();
  const int grid = std::max(1, std::min(rb.num_blocks, 3));  // (grid-stride over work blocks: any grid gives the same result)
  const int* idx = s.idx; double* val = s.val;
  cuda_emu::Launch(grid, kBlock, [=]() { scale_matrix_kernel(rb, idx, val, idx_scale, row_scale); });
}
}  // namespace

// In place: both matrix copies' values, c, q (nullable), the four bound vectors; row_scale / col_scale
// start at 1 and receive the cumulative scaling.
extern "C" int emu_rescale(int n, int m, const int* csr_ptr, const int* csr_idx, double* csr_val, const int* csr_blocks, int csr_num_blocks, const int* csc_ptr,
                           const int* csc_idx, double* csc_val, const int* csc_blocks, int csc_num_blocks, double* c, double* q, double* vlb, double* vub,
                           double* clb, double* cub, int l_inf_ruiz_iterations, int l2_norm_rescaling, int max_grid, double* row, double* col) {
  cuda_emu::defer_async = false;
  HostRows by_row{csr_ptr, csr_idx, csr_val, csr_blocks, csr_num_blocks, m, {}};
  HostRows by_col{csc_ptr, csc_idx, csc_val, csc_blocks, csc_num_blocks, n, {}};
  const RowBlocks rows = by_row.View(), cols = by_col.View();
  std::vector<double> col_norm(std::max(n, 1)), row_norm(std::max(m, 1));
  auto iteration = [&](int l2) {
    Rows(RowNormOp{{}, {}, csc_val, csc_idx, row, col, col_norm.data(), l2}, cols, max_grid);
    Rows(RowNormOp{{}, {}, csr_val, csr_idx, col, row, row_norm.data(), l2}, rows, max_grid);
    Map(DivSqrtOp{{}, {}, col_norm.data(), col}, n);
    Map(DivSqrtOp{{}, {}, row_norm.data(), row}, m);
  };
  bool any = false;
  for (int i = 0; i < l_inf_ruiz_iterations; ++i) { iteration(0); any = true; }
  if (l2_norm_rescaling) { iteration(1); any = true; }
  if (any) {
    Map(ScalePrimalOp{{}, {}, c, vlb, vub, q, col}, n);
    Map(ScaleDualOp{{}, {}, clb, cub, row}, m);
    ScaleMatrix(by_col, row, col);
    ScaleMatrix(by_row, col, row);
  }
  return 0;
}
