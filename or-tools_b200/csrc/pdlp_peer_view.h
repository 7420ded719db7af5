// PeerView: where the slots and flags of the peer-memory exchange live inside a rank's region (plain
// data; shared by the host set-up in pdlp_comm.h and the device code in pdlp_peer.cuh).
#ifndef PDLP_B200_PEER_VIEW_H_
#define PDLP_B200_PEER_VIEW_H_

#include <cstddef>

#if defined(__CUDACC__) && !defined(PDLP_EMULATE)
#define PDLP_HD __host__ __device__
#else
#define PDLP_HD
#endif

namespace pdlp_b200 {

constexpr int kMaxPeers = 8;         // one NVSwitch domain
constexpr int kPeerSumsPerRank = 16;  // doubles per rank in a scalar exchange

// Layout of a region (identical on every rank):
//   [0, 256)                 row flags:  unsigned long long [kMaxPeers] (+ padding)
//   [256, 512)               sums flags: unsigned long long [kMaxPeers]
//   [512, 512 + S)           sums slots: double [2][kMaxPeers][kPeerSumsPerRank], S = 2048
//   [3072, 3076)             own use: how many pushing CTAs of the running launch are done (pdlp_peer.cuh:PeerSharedCtaDone)
//   [4096, ...)              row slots:  double [2][world][cap_rows]
struct PeerView {
  int rank = 0, world = 1;
  long long cap_rows = 0;
  char* base[kMaxPeers] = {};  // base[r] = rank r's region as mapped in this process (base[rank] = own)

  PDLP_HD static constexpr size_t RowSlotsOffset() { return 4096; }
  PDLP_HD static size_t Bytes(int world, long long cap_rows) {
    return RowSlotsOffset() + sizeof(double) * 2 * static_cast<size_t>(world) * static_cast<size_t>(cap_rows);
  }
  PDLP_HD unsigned long long* RowFlags(int owner) const { return reinterpret_cast<unsigned long long*>(base[owner]); }
  PDLP_HD unsigned int* DoneCounter() const { return reinterpret_cast<unsigned int*>(base[rank] + 3072); }
  PDLP_HD unsigned long long* SumFlags(int owner) const { return reinterpret_cast<unsigned long long*>(base[owner] + 256); }
  PDLP_HD double* SumSlot(int owner, int parity, int source) const {
    return reinterpret_cast<double*>(base[owner] + 512) + (static_cast<size_t>(parity) * kMaxPeers + source) * kPeerSumsPerRank;
  }
  PDLP_HD double* RowSlot(int owner, int parity, int source) const {
    return reinterpret_cast<double*>(base[owner] + RowSlotsOffset()) + (static_cast<size_t>(parity) * world + source) * static_cast<size_t>(cap_rows);
  }
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_PEER_VIEW_H_
