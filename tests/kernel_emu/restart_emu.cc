// CPU run of the localized Lagrangian bounds (or-tools_b200/csrc/pdlp_restart_kernels.cuh: the ops AND
// the host loop that finds the trust-region step, the very code the device path runs) with a Runner
// that launches through the shim in cuda_emu.h.  For tests/test_kernel_emulation.py.
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_restart_kernels.cuh"

namespace {
using namespace pdlp_b200;

// K^T y: the plain product op (pdlp_ops.cu RowDotOp's shape).
struct DotOp : AlwaysActive, NoSlots {
  const double* val; const int* idx; const double* v; double* out;
  double Load(int64_t k) const { return val[k] * v[idx[k]]; }
  double RowInit() const { return 0.0; }
  double RowAdd(double acc, double staged, int) const { return acc + staged; }
  double RowMerge(double a, double b) const { return a + b; }
  void RowFinish(int i, double acc, Slots<0>&) const { out[i] = acc; }
};

struct EmuRunner {
  RowBlocks csr, csc;
  const double* csc_val; const int* csc_idx;
  int max_grid;
  std::vector<double> partials = std::vector<double>(static_cast<size_t>(kMaxGrid) * kMaxSlots, 0.0);
  std::vector<double> totals = std::vector<double>(kMaxSlots, 0.0);
  int launches = 0;

  template <class Op>
  const double* Finish(int grid) {
    double* p = partials.data();
    double* t = totals.data();
    if constexpr (Op::kNumSlots > 0) cuda_emu::Launch(1, kBlock, [=]() { finalize_kernel<Op>(p, grid, t); });
    return t;
  }
  template <class Op>
  const double* Rows(const Op& op, const RowBlocks& rb) {
    const int grid = std::max(1, std::min(rb.num_blocks, max_grid));
    double* p = partials.data();
    cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<Op>(op, rb, p); });
    ++launches;
    return Finish<Op>(grid);
  }
  template <class Op>
  const double* CsrRows(const Op& op) { return Rows(op, csr); }
  void ColumnDot(const double* y, double* out) { Rows(DotOp{{}, {}, csc_val, csc_idx, y, out}, csc); }
  template <class Op>
  const double* Map(const Op& op, int64_t count) {
    const int grid = GridFor(count);
    double* p = partials.data();
    cuda_emu::Launch(grid, kBlock, [=]() { map_reduce_kernel<Op>(op, count, p); });
    ++launches;
    return Finish<Op>(grid);
  }
  template <class Op>
  const double* Primal(const Op& op, int64_t count) { return Map(op, count); }
  template <class Op>
  const double* Joint(const Op& op, int64_t count) { return Map(op, count); }
  template <class Op>
  const double* Dual(const Op& op, int64_t count) { return Map(op, count); }
};
}  // namespace

// out = {lagrangian_value, lower_bound, upper_bound, radius, trust-region passes}.
static int g_max_norm = 0;
// The next emu_localized_bounds call computes the max-norm bounds (trust_region.cc:797-884) when set.
extern "C" void emu_set_max_norm(int on) { g_max_norm = on; }
// The next emu_localized_bounds calls take their radius from the distance to this point (null, null: from the argument).
static const double* g_x_start = nullptr;
static const double* g_y_start = nullptr;
// The next calls hand the bounds a precomputed K x (the host-event path: one multi-point pass computes it, pdlp_kkt.cu).
static int g_precomputed_kx = 0;
extern "C" void emu_set_precomputed_kx(int on) { g_precomputed_kx = on; }
extern "C" void emu_set_start_point(const double* x_start, const double* y_start) { g_x_start = x_start; g_y_start = y_start; }

extern "C" int emu_localized_bounds(int n, int m, const int* csr_ptr, const int* csr_idx, const double* csr_val, const int* csr_blocks, int csr_num_blocks,
                                    const int* csc_ptr, const int* csc_idx, const double* csc_val, const int* csc_blocks, int csc_num_blocks, const double* c,
                                    const double* q, const double* vlb, const double* vub, const double* clb, const double* cub, const double* x, const double* y,
                                    double primal_weight, double radius, int use_diagonal_qp_solver, double tolerance, int max_grid, double* out) {
  cuda_emu::defer_async = false;
  std::vector<int> csr_nz(csr_num_blocks + 1), csc_nz(csc_num_blocks + 1);
  for (int b = 0; b <= csr_num_blocks; ++b) csr_nz[b] = csr_ptr[csr_blocks[b]];
  for (int b = 0; b <= csc_num_blocks; ++b) csc_nz[b] = csc_ptr[csc_blocks[b]];
  EmuRunner run{RowBlocks{csr_ptr, csr_blocks, csr_nz.data(), csr_num_blocks, m}, RowBlocks{csc_ptr, csc_blocks, csc_nz.data(), csc_num_blocks, n}, csc_val, csc_idx,
                max_grid};
  std::vector<double> gx(std::max(n, 1)), gy(std::max(m, 1)), kty(std::max(n, 1)), crit(n + m + 1), coef(n + m + 1);
  const RestartProblemView view{csr_val, csr_idx, c, q, vlb, vub, clb, cub, n, m, q == nullptr, 0, gx.data(), gy.data(), kty.data(), crit.data(), coef.data()};
  int passes = 0;
  std::vector<double> kx(std::max(m, 1), 0.0);
  if (g_precomputed_kx != 0) {
    for (int i = 0; i < m; ++i) {
      double a = 0.0;
      for (int k = csr_ptr[i]; k < csr_ptr[i + 1]; ++k) a += csr_val[k] * x[csr_idx[k]];
      kx[i] = a;
    }
  }
  const int launches_before = run.launches;
  const LocalizedBounds b = LocalizedLagrangianBoundsWith(run, view, x, y, nullptr, primal_weight, radius, use_diagonal_qp_solver != 0, tolerance, &passes, g_max_norm != 0, g_x_start, g_y_start,
                                                          g_precomputed_kx != 0 ? kx.data() : nullptr);
  (void)launches_before;
  out[0] = b.lagrangian_value; out[1] = b.lower_bound; out[2] = b.upper_bound; out[3] = b.radius; out[4] = passes;
  return 0;
}
