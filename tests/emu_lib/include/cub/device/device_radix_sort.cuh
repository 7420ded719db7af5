// TEST INFRASTRUCTURE ONLY (see ../../cuda_runtime.h): the one cub entry point the matrix build uses, as a stable
// host sort on the key bits [begin_bit, end_bit) - the contract of cub::DeviceRadixSort::SortPairs.
#ifndef PDLP_TESTS_EMU_CUB_RADIX_SORT_H_
#define PDLP_TESTS_EMU_CUB_RADIX_SORT_H_
#include <algorithm>
#include <numeric>
#include <vector>
#include "cuda_runtime.h"
namespace cub {
struct DeviceRadixSort {
  template <class K, class V>
  static cudaError_t SortPairs(void* temp, size_t& temp_bytes, const K* keys_in, K* keys_out, const V* values_in, V* values_out, int n, int begin_bit, int end_bit,
                               cudaStream_t = nullptr) {
    if (temp == nullptr) { temp_bytes = 1; return cudaSuccess; }
    const unsigned long long mask = end_bit - begin_bit >= 64 ? ~0ull : ((1ull << (end_bit - begin_bit)) - 1ull);
    auto key = [&](int i) { return (static_cast<unsigned long long>(keys_in[i]) >> begin_bit) & mask; };
    std::vector<int> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return key(a) < key(b); });
    for (int i = 0; i < n; ++i) { keys_out[i] = keys_in[order[i]]; values_out[i] = values_in[order[i]]; }
    return cudaSuccess;
  }
};
}  // namespace cub
#endif
