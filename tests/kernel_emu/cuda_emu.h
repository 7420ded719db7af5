// Minimal host stand-in for the CUDA constructs pdlp_skeletons.cuh uses, so that the kernel
// skeletons compile as plain C++ (g++ -std=c++20 -DPDLP_EMULATE) and run on CPU: one CTA at a time,
// one host thread per CUDA thread, __syncthreads and warp shuffles as barriers.  Test
// infrastructure only (tests/test_kernel_emulation.py); it checks index logic and synchronisation
// placement, not performance and not the hardware's memory model.
//
// cp.async is modelled in the two extreme legal orders: copies land at issue ("eager": exposes a
// destination that is overwritten while still needed) or only at wait_group ("deferred": exposes
// data consumed before the wait).
#ifndef PDLP_TESTS_CUDA_EMU_H_
#define PDLP_TESTS_CUDA_EMU_H_

#include <algorithm>
#include <barrier>
#include <bit>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static

namespace cuda_emu {

struct Dim { unsigned x = 0, y = 0, z = 0; };

struct Cta {
  explicit Cta(unsigned threads) : sync(threads), shuffle(threads) {
    for (unsigned w = 0; w < (threads + 31) / 32; ++w) warp_sync.push_back(std::make_unique<std::barrier<>>(std::min(32u, threads - 32 * w)));
  }
  std::barrier<> sync;
  std::vector<std::unique_ptr<std::barrier<>>> warp_sync;
  std::vector<uint64_t> shuffle;
};

inline thread_local Dim thread_idx, block_idx, grid_dim, block_dim;
inline Cta* current_cta = nullptr;
inline bool defer_async = false;  // cp.async lands at wait_group instead of at issue
inline int fdividef_ulps = 0;     // error injected into __fdividef (the intrinsic is good to 2 ulp)

struct PendingCopy { void* dst; const void* src; size_t bytes; };
inline thread_local std::vector<PendingCopy> pending;
inline void AsyncCopy(void* dst, const void* src, size_t bytes) {
  if (defer_async) pending.push_back({dst, src, bytes});
  else std::memcpy(dst, src, bytes);
}
inline void AsyncWaitAll() {
  for (const PendingCopy& p : pending) std::memcpy(p.dst, p.src, p.bytes);
  pending.clear();
}

// Runs body() as a grid of `grid` CTAs of `block` threads, CTAs one after the other.
template <class F>
void Launch(unsigned grid, unsigned block, F body) {
  for (unsigned b = 0; b < grid; ++b) {
    Cta cta(block);
    current_cta = &cta;
    std::vector<std::thread> threads;
    threads.reserve(block);
    for (unsigned t = 0; t < block; ++t) {
      threads.emplace_back([=]() {
        thread_idx = {t, 0, 0}; block_idx = {b, 0, 0}; grid_dim = {grid, 1, 1}; block_dim = {block, 1, 1};
        pending.clear();
        body();
      });
    }
    for (std::thread& th : threads) th.join();
    current_cta = nullptr;
  }
}

}  // namespace cuda_emu

#define threadIdx cuda_emu::thread_idx
#define blockIdx cuda_emu::block_idx
#define gridDim cuda_emu::grid_dim
#define blockDim cuda_emu::block_dim

inline void __syncthreads() { cuda_emu::current_cta->sync.arrive_and_wait(); }
// Every lane of the warp must call (true of every shuffle in the skeletons).
inline double __shfl_xor_sync(unsigned, double v, int lane_mask) {
  cuda_emu::Cta& cta = *cuda_emu::current_cta;
  const unsigned t = cuda_emu::thread_idx.x, w = t >> 5;
  std::memcpy(&cta.shuffle[t], &v, sizeof v);
  cta.warp_sync[w]->arrive_and_wait();
  double r;
  std::memcpy(&r, &cta.shuffle[t ^ static_cast<unsigned>(lane_mask)], sizeof r);
  cta.warp_sync[w]->arrive_and_wait();
  return r;
}
inline int __clz(int x) { return std::countl_zero(static_cast<unsigned>(x)); }
inline int __float2int_rn(float x) { return static_cast<int>(std::nearbyintf(x)); }
inline float __fdividef(float a, float b) {
  float q = a / b;
  for (int i = 0; i < std::abs(cuda_emu::fdividef_ulps); ++i) q = std::nextafterf(q, cuda_emu::fdividef_ulps > 0 ? INFINITY : -INFINITY);
  return q;
}
using std::isfinite;
using std::isinf;
using std::isnan;
inline int min(int a, int b) { return a < b ? a : b; }
inline int max(int a, int b) { return a > b ? a : b; }

#endif  // PDLP_TESTS_CUDA_EMU_H_
