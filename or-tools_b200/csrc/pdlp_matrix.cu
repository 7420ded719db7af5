// Upload of CSC(K) and device-side construction of CSR(K).  One-time setup per solve:
// replaces `transposed_constraint_matrix_ = qp_.constraint_matrix.transpose()`
// (ortools/pdlp/sharded_quadratic_program.cc:66-67).  The stable radix sort is CUB's
// (library code, setup only); everything on the iteration path is hand-written.
#include "pdlp_matrix.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstdlib>
#include <future>
#include <vector>

namespace pdlp_b200 {

namespace {

template <class I>
__global__ void narrow_offsets_kernel(const I* in, int* out, int64_t count, int64_t nnz, int* bad) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
    const int64_t v = in[i];
    if (v < 0 || v > nnz || (i > 0 && in[i - 1] > v) || (i == 0 && v != 0) || (i == count - 1 && v != nnz)) *bad = 1;
    out[i] = static_cast<int>(v);
  }
}

template <class I>
__global__ void narrow_indices_kernel(const I* in, int* out, int* iota, int64_t count, int64_t limit, int* bad) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
    const int64_t v = in[i];
    if (v < 0 || v >= limit) *bad = 2;
    out[i] = static_cast<int>(v);
    iota[i] = static_cast<int>(i);
  }
}

// For each transposed position p: source nonzero perm[p]; its column is the last j with
// col_ptr[j] <= perm[p].
__global__ void gather_transposed_kernel(const int* perm, const int* col_ptr, int n, const double* val, int64_t nnz, int* out_idx, double* out_val) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t p = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < nnz; p += stride) {
    const int src = perm[p];
    int lo = 0, hi = n;  // invariant: col_ptr[lo] <= src < col_ptr[hi]
    while (hi - lo > 1) {
      const int mid = lo + ((hi - lo) >> 1);
      if (col_ptr[mid] <= src) lo = mid; else hi = mid;
    }
    out_idx[p] = lo;
    out_val[p] = val[src];
  }
}

// row_ptr[r] = number of sorted keys < r.
__global__ void row_offsets_kernel(const int* sorted_rows, int64_t nnz, int m, int* row_ptr) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r <= m; r += stride) {
    int64_t lo = 0, hi = nnz;  // first position with key >= r
    while (lo < hi) {
      const int64_t mid = lo + ((hi - lo) >> 1);
      if (sorted_rows[mid] < r) lo = mid + 1; else hi = mid;
    }
    row_ptr[r] = static_cast<int>(lo);
  }
}

// Work-block partition of one orientation, computed on the host from its offsets.
struct HostBlocks {
  std::vector<int> starts, first_nz;
  int max_row_nnz = 0;
  int cut_block = -1;
};

// `cut` in (0, rows): no work block straddles that row; out.cut_block = the block starting there.
template <class P>
HostBlocks ComputeBlocks(const P* ptr, int rows, int cut = -1) {
  HostBlocks out;
  out.starts.push_back(0);
  int mx = 0;
  int r = 0;
  while (r < rows) {
    const int begin = r;
    if (r == cut) out.cut_block = static_cast<int>(out.starts.size()) - 1;
    const int limit = (cut > r) ? cut : rows;
    int64_t nz = static_cast<int64_t>(ptr[r + 1]) - static_cast<int64_t>(ptr[r]);
    mx = nz > mx ? static_cast<int>(nz) : mx;
    ++r;
    if (nz <= kStreamChunk) {
      while (r < limit && r - begin < kMaxRowsPerBlock) {
        const int64_t next = static_cast<int64_t>(ptr[r + 1]) - static_cast<int64_t>(ptr[r]);
        if (nz + next > kStreamChunk) break;
        nz += next;
        mx = next > mx ? static_cast<int>(next) : mx;
        ++r;
      }
    }
    out.starts.push_back(r);
  }
  out.max_row_nnz = mx;
  out.first_nz.resize(out.starts.size());
  for (size_t b = 0; b < out.starts.size(); ++b) out.first_nz[b] = static_cast<int>(ptr[out.starts[b]]);
  return out;
}

void UploadBlocks(SparseRows* s, const HostBlocks& hb, cudaStream_t stream) {
  s->max_row_nnz = hb.max_row_nnz;
  s->num_blocks = static_cast<int>(hb.starts.size()) - 1;
  s->block_start.Resize(hb.starts.size());
  s->block_start.Upload(hb.starts.data(), hb.starts.size(), stream);
  s->block_nz.Resize(hb.first_nz.size());
  s->block_nz.Upload(hb.first_nz.data(), hb.first_nz.size(), stream);
  PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
}

}  // namespace

std::vector<int> PartitionRows(const int* ptr, int rows, int* max_row_nnz) {
  HostBlocks hb = ComputeBlocks(ptr, rows);
  if (max_row_nnz != nullptr) *max_row_nnz = hb.max_row_nnz;
  return std::move(hb.starts);
}

std::vector<int> PartitionRowsWithCut(const int64_t* ptr, int rows, int cut, int* max_row_nnz, int* cut_block) {
  HostBlocks hb = ComputeBlocks(ptr, rows, cut);
  if (max_row_nnz != nullptr) *max_row_nnz = hb.max_row_nnz;
  if (cut_block != nullptr) *cut_block = hb.cut_block;
  return std::move(hb.starts);
}

namespace {
template <class I>
void BuildDeviceMatrixT(int64_t n, int64_t m, int64_t nnz, const I* col_ptr, const I* row_idx, const double* val, cudaStream_t stream,
                        DeviceMatrix* out, int64_t row_cut) {
  if (nnz >= (int64_t{1} << 31) - 1 || n >= (int64_t{1} << 31) - 1 || m >= (int64_t{1} << 31) - 1) {
    throw std::runtime_error("pdlp_b200: one device block must have fewer than 2^31 rows, columns and nonzeros");
  }
  SparseRows& csc = out->by_col;
  SparseRows& csr = out->by_row;
  csc.rows = static_cast<int>(n); csc.cols = static_cast<int>(m); csc.nnz = nnz;
  csr.rows = static_cast<int>(m); csr.cols = static_cast<int>(n); csr.nnz = nnz;
  csc.ptr.Resize(n + 1); csc.idx.Resize(nnz); csc.val.Resize(nnz);
  csr.ptr.Resize(m + 1); csr.idx.Resize(nnz); csr.val.Resize(nnz);

  // The CSC work-block partition only needs the host offsets: it runs on a host thread while the
  // matrix is uploaded and transposed (a sequential scan of n offsets, ~0.1 s at n = 3.3e7).  Offsets
  // that fail the device-side validity check below never reach a kernel: the partition only reads
  // ptr[0 .. n] and its result is dropped with the exception.
  const bool overlap_partition = std::getenv("PDLP_B200_SYNC_PARTITION") == nullptr;  // (timing experiments)
  std::future<HostBlocks> csc_blocks = std::async(overlap_partition ? std::launch::async : std::launch::deferred,
                                                  [col_ptr, n]() { return ComputeBlocks(col_ptr, static_cast<int>(n)); });

  DeviceArray<int> bad(1);
  bad.Zero(stream);
  DeviceArray<int> iota(nnz);
  {
    DeviceArray<I> wide_ptr(n + 1), wide_idx(nnz);
    wide_ptr.Upload(col_ptr, n + 1, stream);
    wide_idx.Upload(row_idx, nnz, stream);
    csc.val.Upload(val, nnz, stream);
    narrow_offsets_kernel<I><<<GridFor(n + 1), kBlock, 0, stream>>>(wide_ptr.get(), csc.ptr.get(), n + 1, nnz, bad.get());
    narrow_indices_kernel<I><<<GridFor(nnz), kBlock, 0, stream>>>(wide_idx.get(), csc.idx.get(), iota.get(), nnz, m, bad.get());
    PDLP_CUDA_CHECK(cudaGetLastError());
    int host_bad = 0;
    bad.Download(&host_bad, 1, stream);
    PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
    if (host_bad != 0) throw std::invalid_argument("pdlp_b200: constraint matrix is not a valid compressed CSC matrix");
  }

  // Stable sort of (row, position) pairs: CSR order with columns ascending per row.
  if (nnz > 0) {
    DeviceArray<int> sorted_rows(nnz), perm(nnz);
    int end_bit = 1;
    while (end_bit < 31 && (int64_t{1} << end_bit) < m) ++end_bit;
    size_t temp_bytes = 0;
    PDLP_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, csc.idx.get(), sorted_rows.get(), iota.get(), perm.get(),
                                                    static_cast<int>(nnz), 0, end_bit, stream));
    DeviceArray<unsigned char> temp(temp_bytes);
    PDLP_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(temp.get(), temp_bytes, csc.idx.get(), sorted_rows.get(), iota.get(), perm.get(),
                                                    static_cast<int>(nnz), 0, end_bit, stream));
    gather_transposed_kernel<<<GridFor(nnz), kBlock, 0, stream>>>(perm.get(), csc.ptr.get(), static_cast<int>(n), csc.val.get(), nnz,
                                                                 csr.idx.get(), csr.val.get());
    row_offsets_kernel<<<GridFor(m + 1), kBlock, 0, stream>>>(sorted_rows.get(), nnz, static_cast<int>(m), csr.ptr.get());
    PDLP_CUDA_CHECK(cudaGetLastError());
    PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
  } else {
    csr.ptr.Zero(stream);
  }

  // Work-block partitions (host, from the offsets).
  UploadBlocks(&csc, csc_blocks.get(), stream);
  std::vector<int> host_ptr(m + 1, 0);
  csr.ptr.Download(host_ptr.data(), m + 1, stream);
  PDLP_CUDA_CHECK(cudaStreamSynchronize(stream));
  const int cut = (row_cut > 0 && row_cut < m) ? static_cast<int>(row_cut) : -1;
  const HostBlocks csr_blocks = ComputeBlocks(host_ptr.data(), static_cast<int>(m), cut);
  UploadBlocks(&csr, csr_blocks, stream);
  // All rows shared (cut at m) or none (cut at 0) are expressed with the block counts alone.
  csr.cut_row = row_cut < 0 ? 0 : static_cast<int>(std::min<int64_t>(row_cut, m));
  csr.cut_block = row_cut < 0 ? -1 : (row_cut >= m ? csr.num_blocks : (row_cut == 0 ? 0 : csr_blocks.cut_block));
  csr.cut_nnz = host_ptr[csr.cut_row];
}
}  // namespace

void BuildDeviceMatrix(int64_t n, int64_t m, int64_t nnz, const int64_t* col_ptr, const int64_t* row_idx, const double* val,
                       cudaStream_t stream, DeviceMatrix* out, int64_t row_cut) {
  BuildDeviceMatrixT(n, m, nnz, col_ptr, row_idx, val, stream, out, row_cut);
}
void BuildDeviceMatrix(int64_t n, int64_t m, int64_t nnz, const int32_t* col_ptr, const int32_t* row_idx, const double* val,
                       cudaStream_t stream, DeviceMatrix* out, int64_t row_cut) {
  BuildDeviceMatrixT(n, m, nnz, col_ptr, row_idx, val, stream, out, row_cut);
}

}  // namespace pdlp_b200
