// Random fp64 gather throughput on one B200: how many nonzeros / s can an SpMV-shaped stream
// (coalesced 4-byte index + coalesced 8-byte value + one 8-byte gather per nonzero) sustain, as a
// function of the load path, the unroll depth and the resident warps.  Upper bound for
// rowstream_kernel.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
// -o gather_bench gather_bench.cu ; run: ./gather_bench [nnz] [n]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

enum Path { kLdg = 0, kNc = 1, kTex = 2, kCg = 3 };

template <int PATH>
__device__ __forceinline__ double Fetch(const double* x, cudaTextureObject_t tex, int j) {
  if constexpr (PATH == kLdg) return x[j];
  else if constexpr (PATH == kNc) return __ldg(x + j);
  else if constexpr (PATH == kCg) return __ldcg(x + j);
  else { int2 v = tex1Dfetch<int2>(tex, j); return __hiloint2double(v.y, v.x); }
}

template <int PATH, int U>
__global__ void __launch_bounds__(256) gather_kernel(const int* __restrict__ idx, const double* __restrict__ val, const double* x,
                                                     cudaTextureObject_t tex, long long nnz, double* out) {
  double acc = 0.0;
  const long long chunk = 256LL * U;
  for (long long base = (long long)blockIdx.x * chunk; base < nnz; base += (long long)gridDim.x * chunk) {
    int j[U]; double v[U], g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { long long k = base + threadIdx.x + u * 256; j[u] = k < nnz ? idx[k] : 0; }
#pragma unroll
    for (int u = 0; u < U; ++u) { long long k = base + threadIdx.x + u * 256; v[u] = k < nnz ? val[k] : 0.0; }
#pragma unroll
    for (int u = 0; u < U; ++u) g[u] = Fetch<PATH>(x, tex, j[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u] * g[u];
  }
  if (acc == 123.456) out[blockIdx.x * 256 + threadIdx.x] = acc;
}

template <int PATH, int U>
float Run(const int* idx, const double* val, const double* x, cudaTextureObject_t tex, long long nnz, double* out, int ctas_per_sm) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  const int grid = 148 * ctas_per_sm;
  for (int w = 0; w < 2; ++w) gather_kernel<PATH, U><<<grid, 256>>>(idx, val, x, tex, nnz, out);
  CK(cudaEventRecord(a));
  const int reps = 5;
  for (int r = 0; r < reps; ++r) gather_kernel<PATH, U><<<grid, 256>>>(idx, val, x, tex, nnz, out);
  CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b));
  return ms / reps;
}

int main(int argc, char** argv) {
  const long long nnz = argc > 1 ? atoll(argv[1]) : 100000000LL;
  const char* pnames[] = {"ldg", "ldg.nc", "tex", "ldcg"};
  for (int pass = 0; pass < 3; ++pass) {
    // pass 0: gathered vector 8 MB (L2-resident); pass 1: 266 MB uniform random (DRAM sectors);
    // pass 2: 266 MB vector, indices random inside a sliding 4 MB window (C3's conservation rows).
    const long long n = pass == 0 ? 1000000LL : 33280000LL;
    std::vector<int> h(nnz);
    std::mt19937_64 rng(1);
    for (long long k = 0; k < nnz; ++k) {
      if (pass < 2) h[k] = (int)(rng() % n);
      else { long long w = 520000, base = (k * (n - w)) / nnz; h[k] = (int)(base + rng() % w); }
    }
    int* idx; double *val, *x, *out;
    CK(cudaMalloc(&idx, nnz * 4)); CK(cudaMalloc(&val, nnz * 8)); CK(cudaMalloc(&x, n * 8)); CK(cudaMalloc(&out, 148 * 16 * 256 * 8));
    CK(cudaMemcpy(idx, h.data(), nnz * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(val, 0, nnz * 8)); CK(cudaMemset(x, 0, n * 8));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = x; rd.res.linear.desc = cudaCreateChannelDesc<int2>(); rd.res.linear.sizeInBytes = n * 8;
    cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex = 0; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    printf("# pass %d: nnz=%lld gathered vector n=%lld (%.0f MB) %s\n", pass, nnz, n, n * 8 / 1e6, pass == 2 ? "windowed" : "uniform");
    printf("path,unroll,ctas_per_sm,ms,Gnnz_per_s,stream_GBps,wavefront_frac_of_1_per_clk\n");
    auto report = [&](int p, int u, int c, float ms) {
      double gn = nnz / (ms * 1e-3) / 1e9;
      printf("%s,%d,%d,%.4f,%.1f,%.0f,%.3f\n", pnames[p], u, c, ms, gn, gn * 12, gn / (148 * 1.965));
    };
    for (int c : {4, 8}) {
      report(0, 1, c, Run<kLdg, 1>(idx, val, x, tex, nnz, out, c));
      report(0, 4, c, Run<kLdg, 4>(idx, val, x, tex, nnz, out, c));
      report(0, 8, c, Run<kLdg, 8>(idx, val, x, tex, nnz, out, c));
      report(0, 16, c, Run<kLdg, 16>(idx, val, x, tex, nnz, out, c));
      report(1, 8, c, Run<kNc, 8>(idx, val, x, tex, nnz, out, c));
      report(3, 8, c, Run<kCg, 8>(idx, val, x, tex, nnz, out, c));
      report(2, 4, c, Run<kTex, 4>(idx, val, x, tex, nnz, out, c));
      report(2, 8, c, Run<kTex, 8>(idx, val, x, tex, nnz, out, c));
    }
    CK(cudaDestroyTextureObject(tex));
    cudaFree(idx); cudaFree(val); cudaFree(x); cudaFree(out);
  }
  return 0;
}
