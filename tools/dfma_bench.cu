// FP64 FMA pipe of one B200 (measurement tool; SURVEY.md section 6 asks for the DFMA peak the fp64 roofline is quoted
// against): dependent-issue latency of DFMA from a single warp with ILP independent chains, and throughput of the whole
// chip as a function of resident warps per SM and ILP.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

template <int ILP>
__global__ void kfma(double* out, int iters, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = threadIdx.x * 1e-3 + i;
  for (int k = 0; k < iters; k++) {
#pragma unroll 16
    for (int u = 0; u < 16; u++)
#pragma unroll
      for (int i = 0; i < ILP; i++) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += x[i];
  if (s == 123.456) out[0] = s;
}

template <int ILP>
static void run(int warps_per_sm, int nsm, double clock_ghz, double* d) {
  const int iters = 2000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int threads = 32 * (warps_per_sm < 4 ? warps_per_sm : 4);
  const int blocks_per_sm = warps_per_sm < 4 ? 1 : warps_per_sm / 4;
  kfma<ILP><<<nsm * blocks_per_sm, threads>>>(d, 10, 1.0000001, 0.5);
  cudaEventRecord(e0);
  kfma<ILP><<<nsm * blocks_per_sm, threads>>>(d, iters, 1.0000001, 0.5);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fmas = (double) nsm * blocks_per_sm * threads * iters * 16.0 * ILP;
  const double per_warp_instr = (double) iters * 16.0 * ILP;
  printf("warps/SM %2d ILP %d: %8.3f ms  %6.2f TFLOP/s  %5.2f cycles per DFMA per warp\n", warps_per_sm, ILP, ms, 2 * fmas / (ms * 1e9),
         ms * 1e-3 * clock_ghz * 1e9 / per_warp_instr);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  printf("%s, %d SMs, %.3f GHz (max)\n", p.name, p.multiProcessorCount, ghz);
  double* d;
  cudaMalloc(&d, 64);
  for (int w: {1, 2, 4, 8, 16, 32}) {
    run<1>(w, p.multiProcessorCount, ghz, d);
    run<2>(w, p.multiProcessorCount, ghz, d);
    run<4>(w, p.multiProcessorCount, ghz, d);
    run<8>(w, p.multiProcessorCount, ghz, d);
  }
  return 0;
}
