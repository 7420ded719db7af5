// TEST INFRASTRUCTURE ONLY.  A host stand-in for the part of the CUDA runtime API that or-tools_b200/csrc uses, so
// that the WHOLE library (host driver + kernels) can be compiled with g++ and run on a machine without a GPU:
// tests/test_full_emulation.py builds tests/_build/libpdlp_b200_emu.so from the same sources (kernel launches rewritten
// by tests/emu_lib/rewrite_launches.py, kernels run by tests/kernel_emu/cuda_emu.h as 256 host threads per CTA) and
// runs device tests against it.  It is never built by __graft_entry__.build(), never installed under or-tools_b200/lib
// and never loaded by the package on its own: the product has no CPU path.
//
// Model: eight identical devices in name (a rank of a sharded solve is a host thread); every stream is synchronous (work runs at the call); "device" memory is host memory; events are
// time stamps.  What this checks is the host driver's logic (launch order, sizes, state hand-over, degenerate
// shapes) together with the kernels' index logic - not timing, not the hardware memory model, not asynchrony.
#ifndef PDLP_TESTS_EMU_CUDA_RUNTIME_H_
#define PDLP_TESTS_EMU_CUDA_RUNTIME_H_

#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "cuda_emu.h"

#define __align__(n) __attribute__((aligned(n)))

typedef int cudaError_t;
enum : int { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2, cudaErrorNotSupported = 801 };
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaMemcpyDefault = 4 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
enum : unsigned { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaHostAllocDefault = 0, cudaHostAllocMapped = 2, cudaIpcMemLazyEnablePeerAccess = 1 };

struct CUstream_st { int id; };
typedef CUstream_st* cudaStream_t;
struct CUevent_st { std::chrono::steady_clock::time_point t; };
typedef CUevent_st* cudaEvent_t;
struct cudaIpcMemHandle_t { char reserved[64]; };

inline const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA runtime: unsupported call or allocation failure"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
// Eight "devices" that differ in name only: a rank of an emulated sharded solve is a host thread of this process.
inline thread_local int emu_current_device = 0;
inline cudaError_t cudaGetDeviceCount(int* n) { *n = 8; return cudaSuccess; }
inline cudaError_t cudaSetDevice(int d) { if (d < 0 || d >= 8) return cudaErrorInvalidValue; emu_current_device = d; return cudaSuccess; }
inline cudaError_t cudaGetDevice(int* d) { *d = emu_current_device; return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr, int) { *v = 2; return cudaSuccess; }  // two "SMs": small grids
template <class K>
inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, K, int, size_t) { *n = 2; return cudaSuccess; }

// Fault injection: the test sets *pdlp_emu_fail_allocation() = k and the k-th cudaMalloc from then on (k >= 1) reports that
// the device is out of memory, once.
extern "C" inline long long* pdlp_emu_fail_allocation() { static long long countdown = 0; return &countdown; }
inline cudaError_t cudaMalloc(void** p, size_t bytes) {
  long long& countdown = *pdlp_emu_fail_allocation();
  if (countdown > 0 && --countdown == 0) { *p = nullptr; return cudaErrorMemoryAllocation; }
  *p = std::aligned_alloc(256, (bytes + 255) / 256 * 256 + 256);
  return *p != nullptr ? cudaSuccess : cudaErrorMemoryAllocation;
}
template <class T> inline cudaError_t cudaMalloc(T** p, size_t bytes) { return cudaMalloc(reinterpret_cast<void**>(p), bytes); }
inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
inline cudaError_t cudaHostAlloc(void** p, size_t bytes, unsigned) { return cudaMalloc(p, bytes); }
inline cudaError_t cudaMallocHost(void** p, size_t bytes) { return cudaMalloc(p, bytes); }
template <class T> inline cudaError_t cudaMallocHost(T** p, size_t bytes) { return cudaMalloc(reinterpret_cast<void**>(p), bytes); }
inline cudaError_t cudaFreeHost(void* p) { std::free(p); return cudaSuccess; }
inline cudaError_t cudaHostGetDevicePointer(void** device, void* host, unsigned) { *device = host; return cudaSuccess; }
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, const T& symbol) { *p = const_cast<void*>(static_cast<const void*>(&symbol)); return cudaSuccess; }

inline cudaError_t cudaMemcpy(void* dst, const void* src, size_t bytes, cudaMemcpyKind) { if (bytes) std::memmove(dst, src, bytes); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* dst, const void* src, size_t bytes, cudaMemcpyKind k, cudaStream_t = nullptr) { return cudaMemcpy(dst, src, bytes, k); }
inline cudaError_t cudaMemset(void* p, int v, size_t bytes) { if (bytes) std::memset(p, v, bytes); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t bytes, cudaStream_t = nullptr) { return cudaMemset(p, v, bytes); }

inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = new CUstream_st{1}; return cudaSuccess; }
inline cudaError_t cudaStreamCreate(cudaStream_t* s) { return cudaStreamCreateWithFlags(s, 0); }
inline cudaError_t cudaStreamDestroy(cudaStream_t s) { delete s; return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }

inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = new CUevent_st{std::chrono::steady_clock::now()}; return cudaSuccess; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { return cudaEventCreateWithFlags(e, 0); }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->t = std::chrono::steady_clock::now(); return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count();
  return cudaSuccess;
}

// Peer memory: the ranks are host threads of one process, so a handle is the pointer itself.
inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t* h, void* p) { std::memset(h, 0, sizeof *h); std::memcpy(h->reserved, &p, sizeof p); h->reserved[63] = 1; return cudaSuccess; }
inline cudaError_t cudaIpcOpenMemHandle(void** p, cudaIpcMemHandle_t h, unsigned) { std::memcpy(p, h.reserved, sizeof *p); return cudaSuccess; }
inline cudaError_t cudaIpcCloseMemHandle(void*) { return cudaSuccess; }

#endif  // PDLP_TESTS_EMU_CUDA_RUNTIME_H_
