// Microbenchmark: cost of straight-line code on sm_100a as a function of its size and of the warps resident per SM.
// Each kernel is N dependent-chain-free DFMAs (8 independent accumulators), fully unrolled: 16 bytes of SASS per DFMA.
// Question it answers: is a 140-260 KB straight-line kernel at 1 warp per scheduler bound by
// instruction fetch?   nvcc -arch=sm_100a -O3 -o icache_bench icache_bench.cu && ./icache_bench
#include <cstdio>
#include <cuda_runtime.h>

template<int N>
__global__ void __launch_bounds__(64) straight(double* out, double c, double d, int reps) {
  extern __shared__ double pad[];
  double a[8];
#pragma unroll
  for (int k = 0; k < 8; k++) a[k] = threadIdx.x + k;
  for (int r = 0; r < reps; r++) {
#pragma unroll
    for (int i = 0; i < N; i++) a[i & 7] = fma(a[i & 7], c, d);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) s += a[k];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s + pad[0];
}

template<int N>
void run(int blocks_per_sm, int smem, int reps) {
  double* out;
  cudaMalloc(&out, 1 << 20);
  cudaFuncSetAttribute(straight<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int grid = 148 * blocks_per_sm * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  straight<N><<<grid, 64, smem>>>(out, 1.0000001, 1e-9, reps);
  cudaEventRecord(e0);
  straight<N><<<grid, 64, smem>>>(out, 1.0000001, 1e-9, reps);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  // 8 waves of blocks_per_sm blocks x 2 warps per SM; cycles per warp-instruction at 1.965 GHz
  const double cyc = ms * 1e-3 * 1.965e9 / (8.0 * reps * N);
  printf("N=%6d (%4d KB code) reps=%d blocks/SM=%d (%d warps/SM): %.3f ms, %.2f cycles per DFMA per warp\n", N, N * 16 / 1024,
         reps, blocks_per_sm, 2 * blocks_per_sm, ms, cyc);
  cudaFree(out);
}

int main() {
  for (int bps: {2, 4}) {
    const int smem = bps == 2 ? 98 * 1024 : 48 * 1024;
    run<512>(bps, smem, 16);
    run<2048>(bps, smem, 4);
    run<8192>(bps, smem, 1);
    run<16384>(bps, smem, 1);
    run<2048>(bps, smem, 1);
    run<512>(bps, smem, 1);
  }
  return 0;
}
