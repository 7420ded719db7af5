// Shared-memory race check of the row-stream kernel: the emulated CTA (cuda_emu.h) run under
// ThreadSanitizer.  Every __syncthreads is a std::barrier, so any pair of conflicting accesses to
// a __shared__ array that is not separated by a barrier (or by the shuffle exchange) is reported as
// a data race.  Exit code 0 = no report.  Built and run by tests/test_kernel_emulation.py:
//   g++ -std=c++20 -O1 -g -fsanitize=thread -pthread race_check.cc
#define PDLP_EMULATE 1
#include "cuda_emu.h"

#include "pdlp_skeletons.cuh"

#include <cstdio>
#include <random>

namespace {
using namespace pdlp_b200;

struct SplitOp {
  static constexpr int kNumSlots = 1;
  static constexpr int SlotKindOf(int) { return kSum; }
  const double* val; const int* idx; const double* x; double* out;
  bool Begin() { return true; }
  const int* IndexAddress(int64_t k) const { return idx + k; }
  const double* ValueAddress(int64_t k) const { return val + k; }
  double Gather(int j) const { return x[j]; }
  double Product(double v, double g) const { return v * g; }
  double Load(int64_t k) const { return val[k] * x[idx[k]]; }
  double RowInit() const { return 0.0; }
  double RowAdd(double acc, double staged, int) const { return acc + staged; }
  double RowMerge(double a, double b) const { return a + b; }
  struct Pre { double o; };
  Pre RowPrefetch(int i) const { return Pre{out[i]}; }
  void RowFinish(int i, double acc, const Pre& pre, Slots<1>& red) const { out[i] = pre.o + acc; red.v[0] += acc; }
};

// Same rule as pdlp_matrix.cu:ComputeBlocks (tests/test_work_blocks.py pins that one).
std::vector<int> Partition(const std::vector<int>& ptr) {
  const int rows = static_cast<int>(ptr.size()) - 1;
  std::vector<int> starts{0};
  int r = 0;
  while (r < rows) {
    const int begin = r;
    long nz = ptr[r + 1] - ptr[r];
    ++r;
    if (nz <= kStreamChunk) {
      while (r < rows && r - begin < kMaxRowsPerBlock && nz + (ptr[r + 1] - ptr[r]) <= kStreamChunk) { nz += ptr[r + 1] - ptr[r]; ++r; }
    }
    starts.push_back(r);
  }
  return starts;
}

int RunShape(const std::vector<int>& lengths, int grid, bool defer) {
  std::vector<int> ptr{0};
  for (int l : lengths) ptr.push_back(ptr.back() + l);
  const int rows = static_cast<int>(lengths.size()), nnz = ptr.back(), cols = 97;
  std::mt19937 gen(7);
  std::vector<int> idx(nnz);
  std::vector<double> val(nnz), x(cols), out(rows, 1.0), want(rows, 1.0);
  for (int& j : idx) j = static_cast<int>(gen() % cols);
  for (double& v : val) v = static_cast<double>(static_cast<int>(gen() % 7) - 3);
  for (double& v : x) v = static_cast<double>(static_cast<int>(gen() % 9) - 4);
  for (int i = 0; i < rows; ++i) for (int k = ptr[i]; k < ptr[i + 1]; ++k) want[i] += val[k] * x[idx[k]];
  const std::vector<int> starts = Partition(ptr);
  const int nb = static_cast<int>(starts.size()) - 1;
  std::vector<int> first(nb + 1);
  for (int b = 0; b <= nb; ++b) first[b] = ptr[starts[b]];
  const RowBlocks rb{ptr.data(), starts.data(), first.data(), nb, rows};
  grid = std::min(grid, nb);
  std::vector<double> partials(grid, 0.0);
  double* p = partials.data();
  const SplitOp op{val.data(), idx.data(), x.data(), out.data()};
  cuda_emu::defer_async = defer;
  cuda_emu::Launch(grid, kBlock, [=]() { rowstream_kernel<SplitOp>(op, rb, p); });
  int bad = 0;
  for (int i = 0; i < rows; ++i) bad += out[i] != want[i];
  return bad;
}
}  // namespace

// The dense skeleton's block reduction, finalize_kernel and the in-kernel partial reduction.
struct SumOp {
  static constexpr int kNumSlots = 2;
  static constexpr int SlotKindOf(int s) { return s == 0 ? kSum : kMax; }
  const double* a;
  bool Begin() { return true; }
  void Apply(int64_t i, Slots<2>& acc) const { acc.v[0] += a[i]; acc.v[1] = fmax(acc.v[1], a[i]); }
};
int RunDense() {
  const int count = 3000, grid = 5;
  std::vector<double> a(count, 1.0), partials(grid * 2, 0.0), totals(3, 0.0);
  a[1234] = 9.0;
  const SumOp op{a.data()};
  double* p = partials.data();
  double* t = totals.data();
  cuda_emu::Launch(grid, kBlock, [=]() { map_reduce_kernel<SumOp>(op, count, p); });
  cuda_emu::Launch(1, kBlock, [=]() { finalize_kernel<SumOp>(p, grid, t); });
  cuda_emu::Launch(1, kBlock, [=]() {
    __shared__ double buf[kBlock];
    const double r = ReducePartialsSum(p, grid, 2, 0, buf);
    const double again = ReducePartialsSum(p, grid, 2, 0, buf);  // back to back, as the decision kernels call it
    if (threadIdx.x == 0) t[2] = r + again;
  });
  return (totals[0] != count + 8.0) + (totals[1] != 9.0) + (totals[2] != 2.0 * (count + 8.0));
}

int main() {
  int bad = RunDense();
  bad += RunShape(std::vector<int>(300, 16), 2, false);                       // full blocks, persistent loop
  std::vector<int> ragged;
  for (int i = 0; i < 500; ++i) ragged.push_back((i * 7) % 23);
  bad += RunShape(ragged, 1, true);                                           // ragged rows, deferred copies
  bad += RunShape({5, 0, 5000, 7, 2048, 2049, 1, 3, 3}, 2, true);             // long rows between staged blocks
  bad += RunShape(std::vector<int>(2500, 1), 1, false);                       // 1024-row blocks, tail loop
  std::printf("wrong rows: %d\n", bad);
  return bad == 0 ? 0 : 1;
}
