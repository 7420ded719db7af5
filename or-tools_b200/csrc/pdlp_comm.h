// The one data-path collective of the column-sharded solve: an
// NCCL all-reduce on the solver's stream.  One process per GPU; rank r owns a contiguous block of
// the columns of K.  NCCL is resolved with dlopen at first use so that the library keeps loading
// on hosts that have no NCCL (the CPU test box) and shares the copy torch already loaded.
#ifndef PDLP_B200_COMM_H_
#define PDLP_B200_COMM_H_

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types and enumerators only; no link dependency

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "pdlp_peer_view.h"

namespace pdlp_b200 {

class NcclApi {
 public:
  static NcclApi& Get() {
    static NcclApi api;
    return api;
  }
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  // Symmetric-memory registration (NCCL >= 2.27); optional: null when the library lacks it.
  ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
  ncclResult_t (*MemFree)(void*) = nullptr;
  ncclResult_t (*CommWindowRegister)(ncclComm_t, void*, size_t, ncclWindow_t*, int) = nullptr;
  ncclResult_t (*CommWindowDeregister)(ncclComm_t, ncclWindow_t) = nullptr;
  bool ok() const { return handle_ != nullptr; }
  bool has_symmetric() const { return MemAlloc != nullptr && MemFree != nullptr && CommWindowRegister != nullptr && CommWindowDeregister != nullptr; }
  const std::string& error() const { return error_; }

 private:
  NcclApi() {
    const char* names[] = {std::getenv("PDLP_B200_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so"};
    for (const char* name : names) {
      if (name == nullptr || name[0] == '\0') continue;
      handle_ = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (handle_ != nullptr) break;
      error_ = dlerror();
    }
    if (handle_ == nullptr) return;
    Load(GetUniqueId, "ncclGetUniqueId");
    Load(CommInitRank, "ncclCommInitRank");
    Load(CommDestroy, "ncclCommDestroy");
    Load(AllReduce, "ncclAllReduce");
    Load(GetErrorString, "ncclGetErrorString");
    if (handle_ != nullptr) {
      MemAlloc = reinterpret_cast<decltype(MemAlloc)>(dlsym(handle_, "ncclMemAlloc"));
      MemFree = reinterpret_cast<decltype(MemFree)>(dlsym(handle_, "ncclMemFree"));
      CommWindowRegister = reinterpret_cast<decltype(CommWindowRegister)>(dlsym(handle_, "ncclCommWindowRegister"));
      CommWindowDeregister = reinterpret_cast<decltype(CommWindowDeregister)>(dlsym(handle_, "ncclCommWindowDeregister"));
    }
  }
  template <class F>
  void Load(F& f, const char* symbol) {
    f = reinterpret_cast<F>(dlsym(handle_, symbol));
    if (f == nullptr) {
      error_ = std::string("missing NCCL symbol ") + symbol;
      handle_ = nullptr;
    }
  }
  void* handle_ = nullptr;
  std::string error_;
};

// Communicators are expensive to create (hundreds of milliseconds), and so is registering an
// exchange buffer with them: both are kept per process, keyed by (unique id, rank, world), and
// reused by every later solve that presents the same id.
struct ExchangeBuffer { void* ptr = nullptr; size_t bytes = 0; ncclWindow_t window = nullptr; bool nccl_memory = false, registered = false; };
// The peer-memory exchange region of this rank and its peers' regions as mapped here (pdlp_peer.cuh).
// Epochs count the exchanges enqueued so far; they continue across solves on the same communicator
// because the flags in the regions do.
struct PeerRegion {
  bool ok = false;
  int rank = 0, world = 1;
  long long cap_rows = 0;
  void* own = nullptr;
  void* mapped[kMaxPeers] = {};  // mapped[rank] == own
  unsigned long long row_epoch = 0, sum_epoch = 0;
};
struct CachedComm {
  ncclComm_t comm = nullptr;
  std::vector<ExchangeBuffer> buffers;  // by slot
  PeerRegion peer;
  int users = 0;                        // live solvers / problem handles on this communicator
};
class CommCache {
 public:
  static CommCache& Get() {
    static CommCache cache;
    return cache;
  }
  CachedComm* Find(const std::string& key) {
    std::lock_guard<std::mutex> lock(mu_);
    auto it = comms_.find(key);
    return it == comms_.end() ? nullptr : &it->second;
  }
  CachedComm* Insert(const std::string& key, ncclComm_t comm) {
    std::lock_guard<std::mutex> lock(mu_);
    CachedComm& c = comms_[key];
    c.comm = comm;
    return &c;
  }
  static void FreeBuffer(ncclComm_t comm, ExchangeBuffer& e) {
    if (e.registered) NcclApi::Get().CommWindowDeregister(comm, e.window);
    if (e.nccl_memory) NcclApi::Get().MemFree(e.ptr);
    else if (e.ptr != nullptr) cudaFree(e.ptr);
    e = ExchangeBuffer();
  }
  // Unmaps the peers' regions, waits until every rank has done so (an exporter must not free memory a
  // peer still has mapped), then frees this rank's region.  Collective.
  static void FreePeerRegion(ncclComm_t comm, PeerRegion& p) {
    if (p.own == nullptr) { p = PeerRegion(); return; }
    cudaDeviceSynchronize();
    for (int r = 0; r < p.world; ++r) {
      if (r != p.rank && p.mapped[r] != nullptr) cudaIpcCloseMemHandle(p.mapped[r]);
    }
    // The first 8 bytes of the own region double as the barrier's buffer (its contents no longer matter).
    NcclApi::Get().AllReduce(p.own, p.own, 1, ncclFloat64, ncclSum, comm, nullptr);
    cudaDeviceSynchronize();
    cudaFree(p.own);
    p = PeerRegion();
  }
  int Release() {  // destroys every cached communicator and its buffers; no solve may be running
    std::lock_guard<std::mutex> lock(mu_);
    const int count = static_cast<int>(comms_.size());
    for (auto& kv : comms_) {
      FreePeerRegion(kv.second.comm, kv.second.peer);
      for (ExchangeBuffer& e : kv.second.buffers) FreeBuffer(kv.second.comm, e);
      NcclApi::Get().CommDestroy(kv.second.comm);
    }
    comms_.clear();
    return count;
  }

 private:
  std::mutex mu_;
  std::map<std::string, CachedComm> comms_;  // node-based: pointers to entries stay valid
};

// rank / world / communicator of one solve.  world == 1: every method is a no-op.
class Comm {
 public:
  Comm() = default;
  Comm(const Comm&) = delete;
  Comm& operator=(const Comm&) = delete;
  ~Comm() { Destroy(); }

  void Init(int rank, int world, const void* unique_id_128) {
    rank_ = rank;
    world_ = world;
    if (world <= 1) return;
    NcclApi& api = NcclApi::Get();
    if (!api.ok()) throw std::runtime_error("pdlp_b200: NCCL is required for a sharded solve but could not be loaded: " + api.error());
    std::string key(static_cast<const char*>(unique_id_128), NCCL_UNIQUE_ID_BYTES);
    key += ":" + std::to_string(rank) + "/" + std::to_string(world);
    cached_ = CommCache::Get().Find(key);
    if (cached_ != nullptr) {
      comm_ = cached_->comm;
      ++cached_->users;
      return;
    }
    ncclUniqueId id;
    std::memcpy(id.internal, unique_id_128, NCCL_UNIQUE_ID_BYTES);
    Check(api.CommInitRank(&comm_, world, id, rank), "ncclCommInitRank");
    cached_ = CommCache::Get().Insert(key, comm_);
    ++cached_->users;
  }
  // Exchange buffer number `slot`, at least `bytes` long (the same request on every rank;
  // collective), for all-reduces in place; kept with the cached communicator and reused by later
  // solves.  PDLP_B200_NCCL_SYMMETRIC=1 allocates it with ncclMemAlloc and registers it as a
  // symmetric window (NCCL >= 2.27) so that NCCL may use its symmetric-memory kernels.  Measured on
  // 8 B200s (C3, 4.2 MB + 32 B per step): 131 vs 138 us for the K x exchange, no change for the
  // 32-byte one, but ~0.27 s of one-time set-up inside the first solve - off by default.
  double* ExchangeBufferFor(int slot, size_t bytes) {
    if (cached_->buffers.size() <= static_cast<size_t>(slot)) cached_->buffers.resize(slot + 1);
    ExchangeBuffer& e = cached_->buffers[slot];
    if (e.ptr != nullptr && e.bytes >= bytes) {
      symmetric_buffers_ += e.registered ? 1 : 0;
      return static_cast<double*>(e.ptr);
    }
    if (e.ptr != nullptr) {
      if (cached_->users > 1) throw std::runtime_error("pdlp_b200: a larger sharded problem cannot be created while another one is alive on the same communicator");
      CommCache::FreeBuffer(comm_, e);
    }
    NcclApi& api = NcclApi::Get();
    const char* env = std::getenv("PDLP_B200_NCCL_SYMMETRIC");
    const bool wanted = api.has_symmetric() && env != nullptr && env[0] == '1';
    if (wanted && api.MemAlloc(&e.ptr, bytes) == ncclSuccess) {
      e.nccl_memory = true;
      e.registered = api.CommWindowRegister(comm_, e.ptr, bytes, &e.window, NCCL_WIN_COLL_SYMMETRIC) == ncclSuccess && e.window != nullptr;
    } else {
      e.ptr = nullptr;
      const cudaError_t err = cudaMalloc(&e.ptr, bytes);
      if (err != cudaSuccess) throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err) + " allocating an exchange buffer");
    }
    e.bytes = bytes;
    cudaMemset(e.ptr, 0, bytes);
    symmetric_buffers_ += e.registered ? 1 : 0;
    return static_cast<double*>(e.ptr);
  }
  int symmetric_buffers() const { return symmetric_buffers_; }

  // Sets up (or reuses) the peer-memory exchange for `cap_rows` shared rows: one region per rank, mapped
  // into every peer with CUDA IPC (the handles travel through one NCCL all-reduce).  Collective: every
  // rank calls it with the same cap_rows.  Returns false - on EVERY rank - when any rank cannot do it
  // (more than kMaxPeers ranks, IPC or peer access unavailable, PDLP_B200_PEER_EXCHANGE=0): the caller
  // then keeps the NCCL all-reduce path.
  bool SetupPeerExchange(long long cap_rows, cudaStream_t stream) {
    if (world_ <= 1) return false;
    PeerRegion& p = cached_->peer;
    const char* env = std::getenv("PDLP_B200_PEER_EXCHANGE");
    bool failed = world_ > kMaxPeers || (env != nullptr && env[0] == '0');
    const bool reuse = !failed && p.ok && p.cap_rows >= cap_rows;
    const bool timing = std::getenv("PDLP_B200_DEBUG_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    double t_alloc = t0, t_handles = t0, t_open = t0;
    if (!failed && !reuse) {
      if (p.own != nullptr && cached_->users > 1) throw std::runtime_error("pdlp_b200: a larger sharded problem cannot be created while another one is alive on the same communicator");
      CommCache::FreePeerRegion(comm_, p);
      p.rank = rank_;
      p.world = world_;
      p.cap_rows = cap_rows;
      const size_t bytes = PeerView::Bytes(world_, cap_rows);
      cudaIpcMemHandle_t handle;
      unsigned char* table = nullptr;  // world x sizeof(handle) bytes: zero except this rank's handle
      const size_t hb = sizeof(cudaIpcMemHandle_t);
      failed = cudaMalloc(&p.own, bytes) != cudaSuccess;
      if (failed) { p.own = nullptr; cudaGetLastError(); }
      if (!failed) failed = cudaMemset(p.own, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess;
      if (!failed) failed = cudaIpcGetMemHandle(&handle, p.own) != cudaSuccess;
      t_alloc = now();
      if (cudaMalloc(&table, hb * world_) != cudaSuccess) throw std::runtime_error("CUDA error: allocating the IPC handle table");
      cudaMemset(table, 0, hb * world_);
      if (!failed) cudaMemcpy(table + hb * rank_, &handle, hb, cudaMemcpyHostToDevice);
      // Bytes summed as uint8: one non-zero contribution per position, so the sum is that rank's byte.
      Check(NcclApi::Get().AllReduce(table, table, hb * world_, ncclUint8, ncclSum, comm_, stream), "ncclAllReduce (IPC handles)");
      std::vector<unsigned char> host(hb * world_);
      cudaMemcpyAsync(host.data(), table, hb * world_, cudaMemcpyDeviceToHost, stream);
      cudaStreamSynchronize(stream);
      cudaFree(table);
      t_handles = now();
      for (int r = 0; r < world_ && !failed; ++r) {
        if (r == rank_) { p.mapped[r] = p.own; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, host.data() + hb * r, hb);
        bool zero = true;
        for (size_t i = 0; i < hb; ++i) zero = zero && host[hb * r + i] == 0;
        if (zero || cudaIpcOpenMemHandle(&p.mapped[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          p.mapped[r] = nullptr;
          cudaGetLastError();
          failed = true;
        }
      }
    }
    t_open = now();
    // Agree: one failure anywhere switches every rank to the NCCL path.
    double flag = failed ? 1.0 : 0.0;
    double* d = ExchangeBufferFor(1, 16 * sizeof(double));
    cudaMemcpyAsync(d, &flag, sizeof(double), cudaMemcpyHostToDevice, stream);
    AllReduce(d, 1, true, stream);
    cudaMemcpyAsync(&flag, d, sizeof(double), cudaMemcpyDeviceToHost, stream);
    cudaStreamSynchronize(stream);
    if (flag != 0.0) {
      CommCache::FreePeerRegion(comm_, p);
      return false;
    }
    p.ok = true;
    if (timing) std::fprintf(stderr, "[pdlp_b200] rank %d peer set-up%s: region %.1f ms, handle exchange %.1f ms, open %.1f ms, agree %.1f ms\n", rank_, reuse ? " (reused)" : "",
                             1e3 * (t_alloc - t0), 1e3 * (t_handles - t_alloc), 1e3 * (t_open - t_handles), 1e3 * (now() - t_open));
    return true;
  }
  bool peer_exchange() const { return cached_ != nullptr && cached_->peer.ok; }
  PeerView peer_view() const {
    PeerView v;
    const PeerRegion& p = cached_->peer;
    v.rank = rank_;
    v.world = world_;
    v.cap_rows = p.cap_rows;
    for (int r = 0; r < world_ && r < kMaxPeers; ++r) v.base[r] = static_cast<char*>(p.mapped[r]);
    return v;
  }
  unsigned long long NextRowEpoch() { ++exchanges_; return ++cached_->peer.row_epoch; }
  unsigned long long NextSumEpoch() { ++exchanges_; return ++cached_->peer.sum_epoch; }
  long long peer_exchanges() const { return exchanges_; }

  void Destroy() {  // the communicator and its exchange buffers stay in the cache
    symmetric_buffers_ = 0;
    if (cached_ != nullptr) --cached_->users;
    cached_ = nullptr;
    comm_ = nullptr;
    world_ = 1;
    rank_ = 0;
  }
  int rank() const { return rank_; }
  int world() const { return world_; }
  bool sharded() const { return world_ > 1; }

  // In-place fp64 all-reduce (sum or max) on `stream`.
  void AllReduce(double* buffer, size_t count, bool max, cudaStream_t stream) {
    if (world_ <= 1 || count == 0) return;
    Check(NcclApi::Get().AllReduce(buffer, buffer, count, ncclFloat64, max ? ncclMax : ncclSum, comm_, stream), "ncclAllReduce");
    ++collectives_;
  }
  long long collectives() const { return collectives_; }

 private:
  void Check(ncclResult_t r, const char* what) {
    if (r != ncclSuccess) throw std::runtime_error(std::string("pdlp_b200: ") + what + " failed: " + NcclApi::Get().GetErrorString(r));
  }
  CachedComm* cached_ = nullptr;
  int symmetric_buffers_ = 0;
  int rank_ = 0, world_ = 1;
  ncclComm_t comm_ = nullptr;
  long long collectives_ = 0;
  long long exchanges_ = 0;  // peer-memory exchanges enqueued (each replaces one all-reduce)
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_COMM_H_
