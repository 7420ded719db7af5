// Minimal host stand-in for the CUDA constructs pdlp_skeletons.cuh uses, so that the kernel
// skeletons compile as plain C++ (g++ -std=c++20 -DPDLP_EMULATE) and run on CPU: one CTA at a time,
// one context per CUDA thread (a host thread each, or with PDLP_EMU_FIBERS user-level contexts switched at barriers by
// the launching thread - what the tests use, except the ThreadSanitizer check), __syncthreads and warp shuffles as barriers.  Test
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
#include <tuple>
#ifdef PDLP_EMU_FIBERS
#include <ucontext.h>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <mutex>
#endif
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#ifdef PDLP_EMU_FIBERS
#define __shared__ static thread_local  // the contexts of a CTA share one host thread; ranks are host threads
#else
#define __shared__ static
#endif

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
#ifdef PDLP_EMU_FIBERS
inline thread_local Cta* current_cta = nullptr;
#else
inline Cta* current_cta = nullptr;
#endif
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

#ifdef PDLP_EMU_FIBERS
// Fiber mode (the whole-library emulation of tests/emu_lib, which makes thousands of launches per solve): the CUDA
// threads of a CTA are 256 user-level contexts run round-robin by the launching host thread; a context runs until its
// next barrier.  A barrier opens when every context that has not finished stands at it: __syncthreads for the whole
// CTA, the barrier inside a shuffle for the warp.  Contexts that wait at barriers which can never open together (half
// a warp at __syncthreads, the other half in a shuffle) are reported as a deadlock - on the device that is a hang.
// The thread mode below stays the default: ThreadSanitizer needs real threads for the race check.
enum class FiberState { kReady, kAtCtaBarrier, kAtWarpBarrier, kDone };
#if defined(__x86_64__) && !defined(PDLP_EMU_UCONTEXT)  // (AddressSanitizer follows swapcontext, not a hand-written switch)
// A context is a stack pointer: the callee-saved registers live on the context's own stack while it is switched out.
// (glibc's swapcontext also saves the signal mask - a system call per switch - which a solve pays millions of times.)
extern "C" void pdlp_emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.weak pdlp_emu_switch
.type pdlp_emu_switch,@function
pdlp_emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size pdlp_emu_switch,.-pdlp_emu_switch
)");
struct EmuContext {
  void* sp = nullptr;
};
inline void EmuSwitch(EmuContext& from, EmuContext& to) { pdlp_emu_switch(&from.sp, to.sp); }
// Prepares `c` to start `entry` (which must never return) on the given stack at its first switch-in.
inline void EmuMakeContext(EmuContext& c, char* stack, size_t bytes, void (*entry)(), EmuContext&) {
  uintptr_t top = (reinterpret_cast<uintptr_t>(stack) + bytes) & ~uintptr_t{15};
  void** sp = reinterpret_cast<void**>(top);
  *--sp = nullptr;                            // the return address `entry` would return to: it does not return
  *--sp = reinterpret_cast<void*>(entry);     // popped by the switch's ret; entry starts with rsp = top - 8
  for (int i = 0; i < 6; ++i) *--sp = nullptr;  // rbp, rbx, r12 - r15
  c.sp = sp;
}
#else
struct EmuContext {
  ucontext_t uc;
};
inline void EmuSwitch(EmuContext& from, EmuContext& to) { swapcontext(&from.uc, &to.uc); }
inline void EmuMakeContext(EmuContext& c, char* stack, size_t bytes, void (*entry)(), EmuContext& back) {
  getcontext(&c.uc);
  c.uc.uc_stack.ss_sp = stack;
  c.uc.uc_stack.ss_size = bytes;
  c.uc.uc_link = &back.uc;
  makecontext(&c.uc, entry, 0);
}
#endif
struct Fiber {
  EmuContext context;
  FiberState state = FiberState::kDone;
  std::vector<PendingCopy> pending;
};
struct FiberScheduler {
  static constexpr size_t kStackBytes = 256 * 1024;
  EmuContext main;
  std::vector<Fiber> fibers;
  std::vector<std::unique_ptr<char[]>> stacks;
  unsigned current = 0;
  const std::function<void()>* body = nullptr;
};
inline thread_local FiberScheduler fiber_scheduler;  // host threads launch independently: `__shared__` is thread_local in this mode

inline void FiberYield(FiberState state) {
  FiberScheduler& s = fiber_scheduler;
  Fiber& f = s.fibers[s.current];
  f.state = state;
  EmuSwitch(f.context, s.main);
}
inline void FiberEntry() {
  FiberScheduler& s = fiber_scheduler;
  (*s.body)();
  for (;;) FiberYield(FiberState::kDone);  // never switched in again
}

template <class F>
void Launch(unsigned grid, unsigned block, F body) {
  FiberScheduler& s = fiber_scheduler;
  if (s.fibers.size() < block) s.fibers.resize(block);
  while (s.stacks.size() < block) s.stacks.push_back(std::make_unique<char[]>(FiberScheduler::kStackBytes));
  const std::function<void()> run = [&]() { body(); };
  s.body = &run;
  grid_dim = {grid, 1, 1};
  block_dim = {block, 1, 1};
  for (unsigned b = 0; b < grid; ++b) {
    Cta cta(block);
    current_cta = &cta;
    block_idx = {b, 0, 0};
    for (unsigned t = 0; t < block; ++t) {
      Fiber& f = s.fibers[t];
      EmuMakeContext(f.context, s.stacks[t].get(), FiberScheduler::kStackBytes, FiberEntry, s.main);
      f.state = FiberState::kReady;
      f.pending.clear();
    }
    unsigned done = 0;
    while (done < block) {
      bool progressed = false;
      for (unsigned t = 0; t < block; ++t) {
        Fiber& f = s.fibers[t];
        if (f.state != FiberState::kReady) continue;
        s.current = t;
        thread_idx = {t, 0, 0};
        pending.swap(f.pending);
        EmuSwitch(s.main, f.context);
        pending.swap(f.pending);
        if (f.state == FiberState::kDone) ++done;
        progressed = true;
      }
      // Open the barriers at which everybody concerned has arrived (finished contexts count as arrived).
      bool cta_open = true;
      for (unsigned t = 0; t < block; ++t) cta_open = cta_open && (s.fibers[t].state == FiberState::kAtCtaBarrier || s.fibers[t].state == FiberState::kDone);
      bool opened = false;
      if (cta_open) {
        for (unsigned t = 0; t < block; ++t) if (s.fibers[t].state == FiberState::kAtCtaBarrier) { s.fibers[t].state = FiberState::kReady; opened = true; }
      }
      for (unsigned w = 0; w * 32 < block; ++w) {
        const unsigned end = std::min(block, w * 32 + 32);
        bool warp_open = true, any = false;
        for (unsigned t = w * 32; t < end; ++t) {
          warp_open = warp_open && (s.fibers[t].state == FiberState::kAtWarpBarrier || s.fibers[t].state == FiberState::kDone);
          any = any || s.fibers[t].state == FiberState::kAtWarpBarrier;
        }
        if (warp_open && any) {
          for (unsigned t = w * 32; t < end; ++t) if (s.fibers[t].state == FiberState::kAtWarpBarrier) s.fibers[t].state = FiberState::kReady;
          opened = true;
        }
      }
      if (!progressed && !opened && done < block) {
        std::fprintf(stderr, "cuda_emu: deadlock in an emulated CTA (block %u of %u): contexts wait at barriers that cannot open together\n", b, grid);
        std::abort();
      }
    }
    current_cta = nullptr;
  }
  s.body = nullptr;
}
#else
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
#endif  // PDLP_EMU_FIBERS

// A kernel launch of the whole-library emulation (tests/emu_lib): the arguments are evaluated once, like the
// parameter buffer of a real launch, and every emulated thread calls the kernel with them.
template <class G, class B, class F, class... A>
void LaunchKernel(G grid, B block, F kernel, A&&... args) {
  const auto parameters = std::make_tuple(std::decay_t<A>(args)...);
  Launch(static_cast<unsigned>(grid), static_cast<unsigned>(block), [&]() { std::apply(kernel, parameters); });
}

}  // namespace cuda_emu

#define threadIdx cuda_emu::thread_idx
#define blockIdx cuda_emu::block_idx
#define gridDim cuda_emu::grid_dim
#define blockDim cuda_emu::block_dim

#ifdef PDLP_EMU_FIBERS
inline void __syncthreads() { cuda_emu::FiberYield(cuda_emu::FiberState::kAtCtaBarrier); }
inline void EmuWarpBarrier(unsigned) { cuda_emu::FiberYield(cuda_emu::FiberState::kAtWarpBarrier); }
#else
inline void __syncthreads() { cuda_emu::current_cta->sync.arrive_and_wait(); }
inline void EmuWarpBarrier(unsigned w) { cuda_emu::current_cta->warp_sync[w]->arrive_and_wait(); }
#endif
// Every lane of the warp must call (true of every shuffle in the skeletons).
inline double __shfl_xor_sync(unsigned, double v, int lane_mask) {
  cuda_emu::Cta& cta = *cuda_emu::current_cta;
  const unsigned t = cuda_emu::thread_idx.x, w = t >> 5;
  std::memcpy(&cta.shuffle[t], &v, sizeof v);
  EmuWarpBarrier(w);
  double r;
  std::memcpy(&r, &cta.shuffle[t ^ static_cast<unsigned>(lane_mask)], sizeof r);
  EmuWarpBarrier(w);
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
