// TEST INFRASTRUCTURE ONLY.  The few NCCL entry points or-tools_b200/csrc/pdlp_comm.h resolves with dlopen, for ranks
// that are host threads of ONE process (tests/test_full_emulation.py: the emulated library's sharded solve).  Built as
// tests/_build/libfake_nccl.so and named through PDLP_B200_NCCL_LIBRARY.  Buffers are host memory; "streams" are
// synchronous, so an all-reduce is a rendezvous of the ranks' threads: everybody publishes its buffer, everybody
// reduces all buffers in rank order into a private copy (bit-identical on every rank, like NCCL's fixed order for a
// fixed world), everybody writes its result back.
#include <algorithm>
#include <barrier>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

extern "C" {
typedef enum { ncclSuccess = 0, ncclInvalidArgument = 4 } ncclResult_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclUint8 = 1, ncclUint64 = 5, ncclFloat64 = 8 } ncclDataType_t;
typedef enum { ncclSum = 0, ncclMax = 2 } ncclRedOp_t;
}

namespace {
struct Group {
  explicit Group(int world) : world(world), rendezvous(world), buffers(world, nullptr) {}
  int world;
  std::barrier<> rendezvous;
  std::vector<const void*> buffers;
};
struct Comm {
  std::shared_ptr<Group> group;
  int rank;
};
std::mutex g_mutex;
std::map<std::string, std::shared_ptr<Group>> g_groups;
uint64_t g_next_id = 1;

template <class T>
void Reduce(const Group& g, T* out, size_t count, bool max) {
  for (size_t i = 0; i < count; ++i) {
    T a = static_cast<const T*>(g.buffers[0])[i];
    for (int r = 1; r < g.world; ++r) {
      const T b = static_cast<const T*>(g.buffers[r])[i];
      a = max ? std::max(a, b) : static_cast<T>(a + b);
    }
    out[i] = a;
  }
}
}  // namespace

extern "C" {
ncclResult_t ncclGetUniqueId(ncclUniqueId* id) {
  std::lock_guard<std::mutex> lock(g_mutex);
  std::memset(id, 0, sizeof *id);
  const uint64_t n = g_next_id++;
  std::memcpy(id->internal, "fake-nccl", 9);
  std::memcpy(id->internal + 16, &n, sizeof n);
  return ncclSuccess;
}
ncclResult_t ncclCommInitRank(Comm** comm, int world, ncclUniqueId id, int rank) {
  if (comm == nullptr || world < 1 || rank < 0 || rank >= world) return ncclInvalidArgument;
  std::shared_ptr<Group> group;
  {
    std::lock_guard<std::mutex> lock(g_mutex);
    std::shared_ptr<Group>& slot = g_groups[std::string(id.internal, sizeof id.internal)];
    if (!slot) slot = std::make_shared<Group>(world);
    if (slot->world != world) return ncclInvalidArgument;
    group = slot;
  }
  *comm = new Comm{group, rank};
  group->rendezvous.arrive_and_wait();  // like the real call, it returns once every rank has joined
  return ncclSuccess;
}
ncclResult_t ncclCommDestroy(Comm* comm) {
  delete comm;
  return ncclSuccess;
}
ncclResult_t ncclAllReduce(const void* send, void* recv, size_t count, ncclDataType_t type, ncclRedOp_t op, Comm* comm, void* /*stream*/) {
  if (comm == nullptr || (op != ncclSum && op != ncclMax)) return ncclInvalidArgument;
  Group& g = *comm->group;
  const size_t width = type == ncclUint8 ? 1 : 8;
  if (type != ncclUint8 && type != ncclUint64 && type != ncclFloat64) return ncclInvalidArgument;
  g.buffers[comm->rank] = send;
  g.rendezvous.arrive_and_wait();
  std::vector<unsigned char> result(count * width);
  if (type == ncclFloat64) Reduce(g, reinterpret_cast<double*>(result.data()), count, op == ncclMax);
  else if (type == ncclUint64) Reduce(g, reinterpret_cast<uint64_t*>(result.data()), count, op == ncclMax);
  else Reduce(g, reinterpret_cast<uint8_t*>(result.data()), count, op == ncclMax);
  g.rendezvous.arrive_and_wait();  // everybody has read every send buffer: in-place results may land now
  if (count > 0) std::memcpy(recv, result.data(), result.size());
  g.rendezvous.arrive_and_wait();
  return ncclSuccess;
}
const char* ncclGetErrorString(ncclResult_t r) { return r == ncclSuccess ? "no error" : "fake NCCL: invalid argument"; }
}  // extern "C"
