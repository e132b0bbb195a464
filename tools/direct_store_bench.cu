// Micro-benchmark (measurement tool, not product code): thread-per-segment output stream written straight from
// registers with 16- or 32-byte vector stores -- no shared-memory transposition -- at a capped number of resident
// warps per SM, against the tile-transposed walk of store_pattern_bench.cu.
//   H<V>: every thread computes V outputs (nf FMAs each; the V chains are independent) and stores them with one
//         st.global.v2.f64 / st.global.v4.f64 at its own region (lane stride = row stride: 32 distinct lines per store).
//   I<V>: the same with the segments of one warp taken `par` apart, so that all lanes share the region's 32-byte phase
//         (row stride 910 doubles: odd segments start 16 bytes off a sector); odd warps skip their first two doubles.
// Usage: direct_store_bench [nseg] [doubles_per_segment] [row_stride] [fma_per_output]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__constant__ int c_stride, c_nf;

template <int V>
__device__ __forceinline__ void vstore(double* p, const double (&v)[V]) {
  if constexpr (V == 4)
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
  else if constexpr (V == 2)
    asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v[0]), "d"(v[1]) : "memory");
  else
    asm volatile("st.global.f64 [%0], %1;" ::"l"(p), "d"(v[0]) : "memory");
}

template <int V, int ILP>
__global__ void __launch_bounds__(32) kH(double* __restrict__ out, int nseg, int dps, int par) {
  // par > 1: warp w handles segments {base + par * lane}: all lanes share (segment mod par)
  const int lane = threadIdx.x;
  int seg;
  int skip = 0;
  if (par > 1) {
    const int group = blockIdx.x / par, ph = blockIdx.x % par;
    seg = group * 32 * par + ph + par * lane;
    skip = (int) (((size_t) seg * c_stride) & 3);  // doubles off a 32-byte boundary
    skip = (4 - skip) & 3;
  } else {
    seg = blockIdx.x * 32 + lane;
  }
  if (seg >= nseg) return;
  double* p = out + (size_t) seg * c_stride + skip;
  double acc[ILP][V];
#pragma unroll
  for (int i = 0; i < ILP; i++)
#pragma unroll
    for (int j = 0; j < V; j++) acc[i][j] = seg * 1e-3 + i + j;
  const int n = (dps - skip) / (V * ILP);
  for (int k = 0; k < n; k++) {
    for (int q = 0; q < c_nf; q++) {
#pragma unroll
      for (int i = 0; i < ILP; i++)
#pragma unroll
        for (int j = 0; j < V; j++) acc[i][j] = fma(acc[i][j], 1.0000001, 0.5);
    }
#pragma unroll
    for (int i = 0; i < ILP; i++) vstore<V>(p + (size_t) (k * ILP + i) * V, acc[i]);
  }
}

int main(int argc, char** argv) {
  const int nseg = argc > 1 ? atoi(argv[1]) : 200000;
  const int dps = argc > 2 ? atoi(argv[2]) : 910;
  const int stride = argc > 3 ? atoi(argv[3]) : dps, nf = argc > 4 ? atoi(argv[4]) : 1;
  const size_t n = (size_t) nseg * dps;
  double* out;
  cudaMalloc(&out, (size_t) nseg * stride * 8 + 4096);
  cudaMemcpyToSymbol(c_stride, &stride, sizeof(int));
  cudaMemcpyToSymbol(c_nf, &nf, sizeof(int));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto timeit = [&](const char* name, auto launch) {
    for (int i = 0; i < 3; i++) launch();
    cudaEventRecord(e0);
    for (int i = 0; i < 10; i++) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= 10;
    printf("%s: %.3f ms  %.0f GB/s  (%s)\n", name, ms, n * 8 / (ms * 1e6), cudaGetErrorString(cudaGetLastError()));
  };
  printf("nseg %d, %d doubles per segment, row stride %d, %d FMA per output, %.1f MB\n", nseg, dps, stride, nf, n * 8 / 1e6);
  const int par = (stride % 4 == 0) ? 1 : (stride % 2 == 0) ? 2 : 4;
#define SETATTR(K) cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)
  SETATTR((kH<1, 4>)); SETATTR((kH<2, 4>)); SETATTR((kH<4, 4>)); SETATTR((kH<4, 1>)); SETATTR((kH<2, 1>)); SETATTR((kH<4, 2>)); SETATTR((kH<2, 8>));
  const int blocks = (nseg + 31) / 32;
  for (int warps: {4, 6, 8, 16, 32}) {
    const size_t dyn = (size_t) (224 * 1024) / warps - 1024;
    char name[96];
    snprintf(name, sizeof(name), "H v1 ilp4 scalar stores,   %2d warps/SM", warps);
    timeit(name, [&] { kH<1, 4><<<blocks, 32, dyn>>>(out, nseg, dps, 1); });
    if (stride % 2 == 0) {
      snprintf(name, sizeof(name), "H v2 ilp1 16-byte stores,  %2d warps/SM", warps);
      timeit(name, [&] { kH<2, 1><<<blocks, 32, dyn>>>(out, nseg, dps, 1); });
      snprintf(name, sizeof(name), "H v2 ilp4 16-byte stores,  %2d warps/SM", warps);
      timeit(name, [&] { kH<2, 4><<<blocks, 32, dyn>>>(out, nseg, dps, 1); });
      snprintf(name, sizeof(name), "H v2 ilp8 16-byte stores,  %2d warps/SM", warps);
      timeit(name, [&] { kH<2, 8><<<blocks, 32, dyn>>>(out, nseg, dps, 1); });
    }
    snprintf(name, sizeof(name), "I v4 ilp1 32-byte, par %d,  %2d warps/SM", par, warps);
    timeit(name, [&] { kH<4, 1><<<blocks, 32, dyn>>>(out, nseg, dps, par); });
    snprintf(name, sizeof(name), "I v4 ilp2 32-byte, par %d,  %2d warps/SM", par, warps);
    timeit(name, [&] { kH<4, 2><<<blocks, 32, dyn>>>(out, nseg, dps, par); });
    snprintf(name, sizeof(name), "I v4 ilp4 32-byte, par %d,  %2d warps/SM", par, warps);
    timeit(name, [&] { kH<4, 4><<<blocks, 32, dyn>>>(out, nseg, dps, par); });
  }
  return 0;
}
