// CPU run of rowstream_kernel<Op> (or-tools_b200/csrc/pdlp_skeletons.cuh) under the shim in
// cuda_emu.h, for tests/test_kernel_emulation.py.  // Every input array is placed right before an inaccessible page, so a read past its end faults.
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_skeletons.cuh"

#include <sys/mman.h>

namespace {
using namespace pdlp_b200;

// K v with an epilogue: split load (indices a block ahead, values by asynchronous copy) and
// epilogue operands prefetched, like the step kernels' ops.
struct SplitOp {
  static constexpr int kNumSlots = 2;
  static constexpr int SlotKindOf(int s) { return s == 0 ? kSum : kMax; }
  const double* val; const int* idx; const double* x; const double* bias; double* out;
  bool Begin() { return true; }
  const int* IndexAddress(int64_t k) const { return idx + k; }
  const double* ValueAddress(int64_t k) const { return val + k; }
  double Gather(int j) const { return x[j]; }
  double Product(double v, double g) const { return v * g; }
  double Load(int64_t k) const { return val[k] * x[idx[k]]; }
  double RowInit() const { return 0.0; }
  double RowAdd(double acc, double staged, int) const { return acc + staged; }
  double RowMerge(double a, double b) const { return a + b; }
  struct Pre { double b, unused0, unused1; };  // > 16 bytes: two prefetched rows per thread, so blocks of many rows take the tail loop
  Pre RowPrefetch(int i) const { return Pre{bias[i], 0.0, 0.0}; }
  void RowFinish(int i, double acc, const Pre& pre, Slots<2>& red) const {
    out[i] = acc + pre.b;
    red.v[0] += acc;
    red.v[1] = fmax(red.v[1], fabs(acc));
  }
};

// Row maxima of |val * x|: plain Load, no prefetch (the scaled-norm ops' shape).
struct PlainMaxOp {
  static constexpr int kNumSlots = 1;
  static constexpr int SlotKindOf(int) { return kMax; }
  const double* val; const int* idx; const double* x; double* out;
  bool Begin() { return true; }
  double Load(int64_t k) const { return fabs(val[k] * x[idx[k]]); }
  double RowInit() const { return 0.0; }
  double RowAdd(double acc, double staged, int) const { return fmax(acc, staged); }
  double RowMerge(double a, double b) const { return fmax(a, b); }
  void RowFinish(int i, double acc, Slots<1>& red) const {
    out[i] = acc;
    red.v[0] = fmax(red.v[0], acc);
  }
};

// Copy of an input array that ends right before an inaccessible page: a read past the end faults.
class Guarded {
 public:
  Guarded(const void* src, size_t bytes) {
    const size_t page = 4096;
    body_ = (bytes + page - 1) / page * page;
    base_ = static_cast<char*>(mmap(nullptr, body_ + page, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0));
    data_ = base_ + body_ - bytes;
    if (bytes > 0) std::memcpy(data_, src, bytes);
    mprotect(base_ + body_, page, PROT_NONE);
  }
  ~Guarded() { munmap(base_, body_ + 4096); }
  template <class T> const T* as() const { return reinterpret_cast<const T*>(data_); }
 private:
  char* base_; char* data_; size_t body_;
};

template <class Op>
void Run(const Op& op, const RowBlocks& rb, int grid, double* totals) {
  std::vector<double> partials(static_cast<size_t>(grid) * Op::kNumSlots, 0.0);
  double* p = partials.data();
  cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<Op>(op, rb, p); });
  for (int s = 0; s < Op::kNumSlots; ++s) {  // finalize_kernel's fixed order
    double t = SlotIdentity(Op::SlotKindOf(s));
    for (int b = 0; b < grid; ++b) t = SlotCombine(Op::SlotKindOf(s), t, partials[static_cast<size_t>(b) * Op::kNumSlots + s]);
    totals[s] = t;
  }
}
}  // namespace

// Dense skeleton: sum, max and a written vector over [0, count), then finalize_kernel and
// ReducePartialsSum over the per-CTA partials.
namespace {
struct AxpyOp {
  static constexpr int kNumSlots = 2;
  static constexpr int SlotKindOf(int s) { return s == 0 ? kSum : kMax; }
  const double* a; double* b; double scale; bool run;
  bool Begin() { return run; }
  void Apply(int64_t i, Slots<2>& acc) const {
    b[i] = b[i] + scale * a[i];
    acc.v[0] += b[i];
    acc.v[1] = fmax(acc.v[1], fabs(b[i]));
  }
};
}  // namespace

// totals[0..1] from finalize_kernel, totals[2] = slot 0 again through ReducePartialsSum.
extern "C" int emu_map_reduce(const double* a, double* b, long long count, double scale, int run, int grid, double* totals) {
  std::vector<double> partials(static_cast<size_t>(grid) * 2, -1.0);
  double* p = partials.data();
  const AxpyOp op{a, b, scale, run != 0};
  cuda_emu::Launch(grid, pdlp_b200::kBlock, [=]() { pdlp_b200::map_reduce_kernel<AxpyOp>(op, count, p); });
  cuda_emu::Launch(1, pdlp_b200::kBlock, [=]() { pdlp_b200::finalize_kernel<AxpyOp>(p, grid, totals); });
  cuda_emu::Launch(1, pdlp_b200::kBlock, [=]() {
    __shared__ double buf[pdlp_b200::kBlock];
    const double r = pdlp_b200::ReducePartialsSum(p, grid, 2, 0, buf);
    if (threadIdx.x == 0) totals[2] = r;
  });
  return 0;
}

extern "C" int emu_stream_chunk() { return pdlp_b200::kStreamChunk; }
extern "C" int emu_max_rows_per_block() { return pdlp_b200::kMaxRowsPerBlock; }

// op_kind 0: SplitOp (out = K x + bias; totals = {sum of K x, max |K x|}); 1: PlainMaxOp.
extern "C" int emu_rowstream(int op_kind, const int* row_ptr, const int* idx, const double* val, const double* x, const double* bias, int rows,
                             const int* block_start, int num_blocks, int grid, int defer_async, int fdividef_ulps, double* out, double* totals) {
  std::vector<int> block_nz(num_blocks + 1);
  for (int b = 0; b <= num_blocks; ++b) block_nz[b] = row_ptr[block_start[b]];
  const size_t nnz = static_cast<size_t>(row_ptr[rows]);
  int cols = 0;
  for (size_t k = 0; k < nnz; ++k) cols = std::max(cols, idx[k] + 1);
  const Guarded g_ptr(row_ptr, (rows + 1) * sizeof(int)), g_idx(idx, nnz * sizeof(int)), g_val(val, nnz * sizeof(double)), g_x(x, cols * sizeof(double)),
      g_bias(bias, rows * sizeof(double)), g_start(block_start, (num_blocks + 1) * sizeof(int)), g_nz(block_nz.data(), (num_blocks + 1) * sizeof(int));
  const pdlp_b200::RowBlocks rb{g_ptr.as<int>(), g_start.as<int>(), g_nz.as<int>(), num_blocks, rows};
  cuda_emu::defer_async = defer_async != 0;
  cuda_emu::fdividef_ulps = fdividef_ulps;
  if (op_kind == 0) Run(SplitOp{g_val.as<double>(), g_idx.as<int>(), g_x.as<double>(), g_bias.as<double>(), out}, rb, grid, totals);
  else if (op_kind == 1) Run(PlainMaxOp{g_val.as<double>(), g_idx.as<int>(), g_x.as<double>(), out}, rb, grid, totals);
  else return 1;
  return 0;
}
