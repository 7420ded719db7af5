// Host <-> device copies of large pageable arrays through pinned staging slots filled by several
// host threads.  A plain cudaMemcpyAsync from pageable memory is staged by the driver on one thread
// (~10 GB/s measured on the C3 upload: 2.7 GB in 0.28 s); here T threads (up to 12) each own two 4 MB pinned slots
// and a stream, copy their chunks into a slot and queue the DMA, so that the host-side copies run in
// parallel and overlap the transfers.  The problem upload and the solution download are the only
// host <-> device traffic of a solve, so this is what the end-to-end time sees of PCIe.
//
// Both calls are synchronous for the host buffer (it may be reused / is filled on return) and
// stream-ordered for the device buffer: the transfer starts after the work already queued on
// `stream` and later work on `stream` waits for it.
#ifndef PDLP_B200_STAGER_H_
#define PDLP_B200_STAGER_H_

#include <cuda_runtime.h>

#include <algorithm>
#include <cstddef>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace pdlp_b200 {

class HostStager {
 public:
  static constexpr size_t kChunkBytes = size_t{4} << 20;
  static constexpr size_t kMinBytes = size_t{4} << 20;  // smaller copies take the plain path
  static constexpr int kSlots = 2;

  static HostStager& Get() {
    static HostStager instance;
    return instance;
  }

  // Returns false when staging is unavailable (pinned allocation failed): caller falls back to cudaMemcpyAsync.
  bool Upload(void* device_dst, const void* host_src, size_t bytes, cudaStream_t stream) { return Run(device_dst, const_cast<void*>(host_src), bytes, stream, true); }
  bool Download(void* host_dst, const void* device_src, size_t bytes, cudaStream_t stream) { return Run(const_cast<void*>(device_src), host_dst, bytes, stream, false); }

 private:
  struct Lane {
    cudaStream_t stream = nullptr;
    void* slot[kSlots] = {nullptr, nullptr};
    cudaEvent_t done[kSlots] = {nullptr, nullptr};
  };

  HostStager() {
    const unsigned hw = std::thread::hardware_concurrency();
    // One host thread per lane; leave two cores to the caller (2 x 4 MB pinned per lane, <= 96 MB).
    const int lanes = static_cast<int>(std::max(1u, std::min(12u, hw > 2 ? hw - 2 : 1u)));
    lanes_.resize(lanes);
  }
  ~HostStager() = default;  // pinned memory and streams live for the process (the context may be gone at exit)

  // Pinned slots belong to a device context: (re)create them for the current device.
  bool Prepare() {
    int device = -1;
    if (cudaGetDevice(&device) != cudaSuccess) return false;
    if (device == device_ && ready_) return true;
    Release();
    for (Lane& lane : lanes_) {
      if (cudaStreamCreateWithFlags(&lane.stream, cudaStreamNonBlocking) != cudaSuccess) return Fail();
      for (int s = 0; s < kSlots; ++s) {
        if (cudaHostAlloc(&lane.slot[s], kChunkBytes, cudaHostAllocDefault) != cudaSuccess) return Fail();
        if (cudaEventCreateWithFlags(&lane.done[s], cudaEventDisableTiming) != cudaSuccess) return Fail();
      }
    }
    if (cudaEventCreateWithFlags(&begin_, cudaEventDisableTiming) != cudaSuccess) return Fail();
    device_ = device;
    ready_ = true;
    return true;
  }
  bool Fail() {
    cudaGetLastError();
    Release();
    return false;
  }
  void Release() {
    for (Lane& lane : lanes_) {
      for (int s = 0; s < kSlots; ++s) {
        if (lane.slot[s] != nullptr) cudaFreeHost(lane.slot[s]);
        if (lane.done[s] != nullptr) cudaEventDestroy(lane.done[s]);
        lane.slot[s] = nullptr;
        lane.done[s] = nullptr;
      }
      if (lane.stream != nullptr) cudaStreamDestroy(lane.stream);
      lane.stream = nullptr;
    }
    if (begin_ != nullptr) cudaEventDestroy(begin_);
    begin_ = nullptr;
    ready_ = false;
    device_ = -1;
  }

  bool Run(void* device_ptr, void* host_ptr, size_t bytes, cudaStream_t stream, bool upload) {
    std::lock_guard<std::mutex> lock(mu_);
    if (!Prepare()) return false;
    const size_t chunks = (bytes + kChunkBytes - 1) / kChunkBytes;
    const int lanes = static_cast<int>(std::min<size_t>(lanes_.size(), chunks));
    // Lane streams start after what is already queued on the caller's stream.
    if (cudaEventRecord(begin_, stream) != cudaSuccess) return false;
    int device = device_;
    std::vector<cudaError_t> status(lanes, cudaSuccess);
    auto work = [&](int t) {
      Lane& lane = lanes_[t];
      cudaError_t err = cudaSetDevice(device);
      if (err == cudaSuccess) err = cudaStreamWaitEvent(lane.stream, begin_, 0);
      char* dev = static_cast<char*>(device_ptr);
      char* host = static_cast<char*>(host_ptr);
      if (upload) {
        int use = 0;
        for (size_t c = t; c < chunks && err == cudaSuccess; c += lanes, use ^= 1) {
          const size_t off = c * kChunkBytes, len = std::min(kChunkBytes, bytes - off);
          err = cudaEventSynchronize(lane.done[use]);  // slot free again (a never-recorded event is complete)
          if (err != cudaSuccess) break;
          std::memcpy(lane.slot[use], host + off, len);
          err = cudaMemcpyAsync(dev + off, lane.slot[use], len, cudaMemcpyHostToDevice, lane.stream);
          if (err == cudaSuccess) err = cudaEventRecord(lane.done[use], lane.stream);
        }
      } else {
        // Two-deep: the next chunk's DMA runs while this chunk is copied out of its slot.
        size_t pending_off = 0, pending_len = 0;
        int pending_slot = -1, use = 0;
        for (size_t c = t; c < chunks && err == cudaSuccess; c += lanes, use ^= 1) {
          const size_t off = c * kChunkBytes, len = std::min(kChunkBytes, bytes - off);
          err = cudaMemcpyAsync(lane.slot[use], dev + off, len, cudaMemcpyDeviceToHost, lane.stream);
          if (err == cudaSuccess) err = cudaEventRecord(lane.done[use], lane.stream);
          if (pending_slot >= 0 && err == cudaSuccess) {
            err = cudaEventSynchronize(lane.done[pending_slot]);
            if (err == cudaSuccess) std::memcpy(host + pending_off, lane.slot[pending_slot], pending_len);
          }
          pending_off = off; pending_len = len; pending_slot = use;
        }
        if (pending_slot >= 0 && err == cudaSuccess) {
          err = cudaEventSynchronize(lane.done[pending_slot]);
          if (err == cudaSuccess) std::memcpy(host + pending_off, lane.slot[pending_slot], pending_len);
        }
      }
      status[t] = err;
    };
    std::vector<std::thread> threads;
    threads.reserve(lanes > 0 ? lanes - 1 : 0);
    for (int t = 1; t < lanes; ++t) threads.emplace_back(work, t);
    work(0);
    for (std::thread& th : threads) th.join();
    bool ok = true;
    for (int t = 0; t < lanes; ++t) ok = ok && status[t] == cudaSuccess;
    // The caller's stream continues after every lane's last transfer.
    for (int t = 0; t < lanes && ok; ++t) {
      for (int s = 0; s < kSlots; ++s) ok = ok && cudaStreamWaitEvent(stream, lanes_[t].done[s], 0) == cudaSuccess;
    }
    if (!ok) {
      cudaGetLastError();
      // Leave no transfer in flight behind a failed call.
      for (Lane& lane : lanes_) cudaStreamSynchronize(lane.stream);
    }
    return ok;
  }

  std::mutex mu_;
  std::vector<Lane> lanes_;
  cudaEvent_t begin_ = nullptr;
  int device_ = -1;
  bool ready_ = false;
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_STAGER_H_
