// The one data-path collective of the column-sharded solve: an
// NCCL all-reduce on the solver's stream.  One process per GPU; rank r owns a contiguous block of
// the columns of K.  NCCL is resolved with dlopen at first use so that the library keeps loading
// on hosts that have no NCCL (the CPU test box) and shares the copy torch already loaded.
#ifndef PDLP_B200_COMM_H_
#define PDLP_B200_COMM_H_

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types and enumerators only; no link dependency

#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>

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
  bool ok() const { return handle_ != nullptr; }
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

// Communicators are expensive to create (hundreds of milliseconds): they are kept per process,
// keyed by (unique id, rank, world), and reused by every later solve that presents the same id.
class CommCache {
 public:
  static CommCache& Get() {
    static CommCache cache;
    return cache;
  }
  ncclComm_t Find(const std::string& key) {
    std::lock_guard<std::mutex> lock(mu_);
    auto it = comms_.find(key);
    return it == comms_.end() ? nullptr : it->second;
  }
  void Insert(const std::string& key, ncclComm_t comm) {
    std::lock_guard<std::mutex> lock(mu_);
    comms_[key] = comm;
  }
  int Release() {  // destroys every cached communicator; no solve may be running
    std::lock_guard<std::mutex> lock(mu_);
    const int count = static_cast<int>(comms_.size());
    for (auto& kv : comms_) NcclApi::Get().CommDestroy(kv.second);
    comms_.clear();
    return count;
  }

 private:
  std::mutex mu_;
  std::map<std::string, ncclComm_t> comms_;
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
    comm_ = CommCache::Get().Find(key);
    if (comm_ != nullptr) return;
    ncclUniqueId id;
    std::memcpy(id.internal, unique_id_128, NCCL_UNIQUE_ID_BYTES);
    Check(api.CommInitRank(&comm_, world, id, rank), "ncclCommInitRank");
    CommCache::Get().Insert(key, comm_);
  }
  void Destroy() {  // the communicator stays in the cache
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
  int rank_ = 0, world_ = 1;
  ncclComm_t comm_ = nullptr;
  long long collectives_ = 0;
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_COMM_H_
