// Device plumbing shared by every PDLP kernel file: error checking, RAII device buffers, and
// (pdlp_skeletons.cuh) deterministic block reductions and the two kernel skeletons every op is an
// instance of:
//
//   map_reduce_kernel<Op>   one grid-stride pass over a dense index range [0, count)
//   rowstream_kernel<Op>    one pass over a compressed sparse matrix by rows ("row" = a
//                           column of the CSC copy or a row of the CSR copy), the device
//                           replacement of TransposedMatrixVectorProduct
//                           (ortools/pdlp/sharder.cc:163-176) and of the scaled column
//                           norms (sharder.cc:291-335).
//
// Both skeletons end with a fixed-shape two-stage reduction (thread -> warp shuffle ->
// shared memory -> one partial per block, then a finalize pass in fixed order), so
// results are bit-reproducible for a given launch configuration: the device analogue of
// the reference's "same num_shards => same roundoff" promise (solvers.proto:273-282).
#ifndef PDLP_B200_COMMON_CUH_
#define PDLP_B200_COMMON_CUH_

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

#include "pdlp_stager.h"

namespace pdlp_b200 {

#define PDLP_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t err__ = (expr);                                                            \
    if (err__ != cudaSuccess) {                                                            \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err__) +  \
                               " at " + __FILE__ + ":" + std::to_string(__LINE__));       \
    }                                                                                      \
  } while (0)

// Owning device array.  Allocation is synchronous (cudaMalloc); all solver buffers are
// created once per solve, never inside the iteration loop.
template <typename T>
class DeviceArray {
 public:
  DeviceArray() = default;
  explicit DeviceArray(size_t n) { Resize(n); }
  DeviceArray(const DeviceArray&) = delete;
  DeviceArray& operator=(const DeviceArray&) = delete;
  DeviceArray(DeviceArray&& o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr; o.n_ = 0; }
  DeviceArray& operator=(DeviceArray&& o) noexcept {
    if (this != &o) { Free(); p_ = o.p_; n_ = o.n_; o.p_ = nullptr; o.n_ = 0; }
    return *this;
  }
  ~DeviceArray() { Free(); }
  void Resize(size_t n) {
    Free();
    n_ = n;
    if (n > 0) PDLP_CUDA_CHECK(cudaMalloc(&p_, n * sizeof(T)));
  }
  void Free() {
    if (p_ != nullptr) cudaFree(p_);
    p_ = nullptr;
    n_ = 0;
  }
  T* get() const { return p_; }
  size_t size() const { return n_; }
  // Large arrays travel through the multi-threaded pinned stager (pdlp_stager.h); small ones and
  // the fallback through a plain stream-ordered copy.
  void Upload(const T* host, size_t n, cudaStream_t s) {
    if (n == 0) return;
    if (n * sizeof(T) >= HostStager::kMinBytes && HostStager::Get().Upload(p_, host, n * sizeof(T), s)) return;
    PDLP_CUDA_CHECK(cudaMemcpyAsync(p_, host, n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void Download(T* host, size_t n, cudaStream_t s) const {
    if (n == 0) return;
    if (n * sizeof(T) >= HostStager::kMinBytes && HostStager::Get().Download(host, p_, n * sizeof(T), s)) return;
    PDLP_CUDA_CHECK(cudaMemcpyAsync(host, p_, n * sizeof(T), cudaMemcpyDeviceToHost, s));
  }
  void Zero(cudaStream_t s) {
    if (n_ > 0) PDLP_CUDA_CHECK(cudaMemsetAsync(p_, 0, n_ * sizeof(T), s));
  }
  void swap(DeviceArray& o) { std::swap(p_, o.p_); std::swap(n_, o.n_); }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};
using DVec = DeviceArray<double>;

}  // namespace pdlp_b200

#include "pdlp_skeletons.cuh"

#endif  // PDLP_B200_COMMON_CUH_
