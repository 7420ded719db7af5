// K v from caller-owned device arrays without any allocation: the kernel side of
// pdlp_b200_matvec_no_alloc (include/pdlp_b200.h).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).
//
// The row-stream kernel (pdlp_skeletons.cuh) needs, besides the matrix, a work-block table: where every
// block of consecutive rows starts and its first nonzero.  pdlp_b200_problem_create computes that table
// on the host and keeps it in cudaMalloc'ed arrays; here it is built ON THE DEVICE into a file-scope
// static array, so the call allocates nothing:
//   1. fits_kernel      for R = 1024, 512, ... 1 rows per block: does any block of R consecutive rows hold
//                       more than kStreamChunk nonzeros?  (R = 1 always fits: a single long row is allowed)
//   2. choose_kernel    the largest R that fits; number of blocks = ceil(rows / R)
//   3. fill_kernel      block_start[b] = min(b R, rows), block_nz[b] = row_ptr[block_start[b]]
//   4. rowstream_kernel<RowDotOp> over that table: the same kernel pdlp_b200_matvec launches.
// The uniform-R table is coarser than the host's greedy one (fewer rows per block when row lengths
// vary), which costs speed, not correctness; threads per row may differ from the greedy table's, so sums
// of rows that several lanes share are reassociated: results agree with pdlp_b200_matvec to rounding.
//
// The scratch array is shared by every call on a device and keeps the last call's contents: the flags of
// step 1 are cleared at the start of each call, everything else is overwritten before it is read.  Two
// calls that overlap on the same GPU (different streams or threads) race on it: callers serialise.
#include <mutex>
#include <stdexcept>
#include <string>

#include "pdlp_b200.h"
#include "pdlp_ops.cuh"
#include "pdlp_ops_kernels.cuh"

namespace pdlp_b200 {

namespace {

constexpr int kNoAllocMaxRows = PDLP_B200_NO_ALLOC_MAX_ROWS;
constexpr int kLevels = 11;  // R = 1024 >> level
static_assert((1024 >> (kLevels - 1)) == 1 && kMaxRowsPerBlock >= 1024, "levels must end at one row per block");

struct NoAllocScratch {
  int violated[kLevels];  // step 1: some block of 1024 >> level rows is too long
  int rows_per_block;     // step 2
  int num_blocks;
  int reserved[3];
  int block_start[kNoAllocMaxRows + 1];
  int block_nz[kNoAllocMaxRows + 1];
};
// ~2 MB, held by every GPU that loads the library until the process exits.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

__global__ void __launch_bounds__(kBlock) fits_kernel(NoAllocScratch* s, const int* row_ptr, int rows) {
  const int64_t g = static_cast<int64_t>(blockIdx.x) * kBlock + threadIdx.x;
  for (int level = 0; level < kLevels - 1; ++level) {  // one row per block needs no test
    const int r = 1024 >> level;
    const int64_t first = g * r;
    if (first >= rows) continue;  // coarser levels have fewer groups; finer levels of this thread start further left
    const int64_t last = first + r < rows ? first + r : rows;
    if (row_ptr[last] - row_ptr[first] > kStreamChunk) s->violated[level] = 1;
  }
}

__global__ void choose_kernel(NoAllocScratch* s, int rows) {
  int level = 0;
  while (level < kLevels - 1 && s->violated[level] != 0) ++level;
  const int r = 1024 >> level;
  s->rows_per_block = r;
  s->num_blocks = (rows + r - 1) / r;
}

__global__ void __launch_bounds__(kBlock) fill_kernel(NoAllocScratch* s, const int* row_ptr, int rows) {
  const int nb = s->num_blocks, r = s->rows_per_block;
  const int stride = gridDim.x * kBlock;
  for (int b = blockIdx.x * kBlock + threadIdx.x; b <= nb; b += stride) {
    const int64_t start = static_cast<int64_t>(b) * r;
    const int row = start < rows ? static_cast<int>(start) : rows;
    s->block_start[b] = row;
    s->block_nz[b] = row_ptr[row];
  }
}

// The device address of the scratch array, looked up once per device with that device current.  In host
// code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel could use.
NoAllocScratch* ScratchOfCurrentDevice() {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static NoAllocScratch* address[kMaxDevices] = {};
  int device = 0;
  PDLP_CUDA_CHECK(cudaGetDevice(&device));
  if (device < 0 || device >= kMaxDevices) throw std::runtime_error("pdlp_b200_matvec_no_alloc: device ordinal beyond 63");
  std::lock_guard<std::mutex> lock(mu);
  if (address[device] == nullptr) {
    void* p = nullptr;
    PDLP_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    address[device] = static_cast<NoAllocScratch*>(p);
  }
  return address[device];
}

}  // namespace

// out[rows] = K v for the CSR arrays (row_ptr: rows + 1 offsets, col_idx / values: row_ptr[rows] entries),
// all device pointers on the current device; enqueued on `stream` and waited for.  Throws
// std::invalid_argument for sizes beyond the scratch array.
void MatVecNoAlloc(int64_t rows, int64_t cols, const int* row_ptr, const int* col_idx, const double* values, const double* v, double* out, cudaStream_t stream) {
  if (rows > kNoAllocMaxRows) {
    throw std::invalid_argument("pdlp_b200_matvec_no_alloc: " + std::to_string(rows) + " rows exceed the static scratch limit of " +
                                std::to_string(kNoAllocMaxRows) + " rows (use pdlp_b200_problem_create + pdlp_b200_matvec)");
  }
  (void)cols;
  if (rows == 0) return;
  NoAllocScratch* s = ScratchOfCurrentDevice();
  const int n = static_cast<int>(rows);
  PDLP_CUDA_CHECK(cudaMemsetAsync(s, 0, sizeof(int) * (kLevels + 5), stream));  // flags and the two scalars: nothing of the last call survives
  const int groups = (n + kBlock - 1) / kBlock;
  fits_kernel<<<groups, kBlock, 0, stream>>>(s, row_ptr, n);
  choose_kernel<<<1, 1, 0, stream>>>(s, n);
  fill_kernel<<<GridFor(n + 1), kBlock, 0, stream>>>(s, row_ptr, n);
  int num_blocks = 0;
  PDLP_CUDA_CHECK(cudaMemcpyAsync(&num_blocks, &s->num_blocks, sizeof(int), cudaMemcpyDeviceToHost, stream));
  PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (num_blocks < 1 || num_blocks > n) throw std::runtime_error("pdlp_b200_matvec_no_alloc: inconsistent work-block table (are row_ptr's offsets ascending?)");
  const RowBlocks rb{row_ptr, s->block_start, s->block_nz, num_blocks, n};
  const RowDotOp op{{}, {}, values, col_idx, v, out};
  const int grid = num_blocks < kMaxGrid / 2 ? num_blocks : kMaxGrid / 2;  // 4 resident CTAs per SM, as RowStreamGrid finds for this kernel
  rowstream_kernel<RowDotOp><<<grid, kBlock, 0, stream>>>(op, rb, nullptr);
  PDLP_CUDA_CHECK(cudaGetLastError());
  PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
}

}  // namespace pdlp_b200
