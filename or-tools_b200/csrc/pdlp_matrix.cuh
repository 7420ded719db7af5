// Device-resident constraint matrix in both orientations, the replacement of
// ShardedQuadraticProgram's `qp_.constraint_matrix` + `transposed_constraint_matrix_`
// (ortools/pdlp/sharded_quadratic_program.cc:63-93).  fp64 values, int32 indices and
// offsets (the reference uses int64, quadratic_program.h:130; one GPU's block must
// hold fewer than 2^31 nonzeros).
#ifndef PDLP_B200_MATRIX_CUH_
#define PDLP_B200_MATRIX_CUH_

#include <vector>

#include "pdlp_common.cuh"

namespace pdlp_b200 {

// A matrix stored by compressed "rows".  For CSC(K) a "row" is a column of K.
struct SparseRows {
  int rows = 0, cols = 0;
  int64_t nnz = 0;
  DeviceArray<int> ptr;  // rows + 1
  DeviceArray<int> idx;  // nnz
  DVec val;              // nnz
  DeviceArray<int> block_start;  // num_blocks + 1
  DeviceArray<int> block_nz;     // num_blocks + 1: ptr[block_start[b]]
  int num_blocks = 0;
  int max_row_nnz = 0;
  // Column-sharded solve: the CSC block's compressed "rows" are this rank's columns (their
  // reductions are per-rank partials); the CSR block holds only this rank's part of every row sum.
  bool sharded_rows = false, partial_sums = false;
  // Column-sharded CSR block: work blocks [0, cut_block) hold exactly the rows shared with other
  // ranks (rows [0, cut_row)), the rest the rows private to this rank.  -1: no cut.
  int cut_block = -1, cut_row = 0;
  int64_t cut_nnz = 0;  // nonzeros of the rows above the cut (the shared part's share of the work)
  RowBlocks View() const { return RowBlocks{ptr.get(), block_start.get(), block_nz.get(), num_blocks, rows}; }
  // Work blocks [first, first + count): same row / nonzero numbering, so ops index rows as usual.
  RowBlocks ViewOfBlocks(int first, int count) const { return RowBlocks{ptr.get(), block_start.get() + first, block_nz.get() + first, count, rows}; }
};

struct DeviceMatrix {
  SparseRows by_col;  // CSC of K: n compressed columns, indices are constraint rows
  SparseRows by_row;  // CSR of K: m compressed rows, indices are variable columns
};

// Cuts [0, rows) into work blocks for rowstream_kernel from a host copy of the offsets.
std::vector<int> PartitionRows(const int* ptr, int rows, int* max_row_nnz);
// Host-only view of the same partition for the unit tests (pdlp_b200_work_blocks): block starts
// with a cut at row `cut` (-1: none); *cut_block = index of the block that starts there, or -1.
std::vector<int> PartitionRowsWithCut(const int64_t* ptr, int rows, int cut, int* max_row_nnz, int* cut_block);

// Uploads CSC(K) given with int64 offsets/indices and builds CSR(K) on the device
// (stable: columns ascending inside every row, exactly the order of Eigen's
// `.transpose()` materialisation).  Throws std::runtime_error on CUDA failure.
// `row_cut` >= 0 forces a work-block boundary of the CSR copy at that row (the shared / private split of
// a column-sharded block).
void BuildDeviceMatrix(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val,
                       cudaStream_t stream, DeviceMatrix* out, int64_t row_cut = -1);
// The same from 32-bit offsets / indices (what scipy holds below 2^31 nonzeros): half the index upload.
void BuildDeviceMatrix(int64_t n, int64_t m, int64_t nnz, const int32_t* col_ptr, const int32_t* row_idx, const double* val,
                       cudaStream_t stream, DeviceMatrix* out, int64_t row_cut = -1);

}  // namespace pdlp_b200
#endif  // PDLP_B200_MATRIX_CUH_
