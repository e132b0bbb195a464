// Micro-benchmark of the output store patterns of the segment kernels (measurement tool, not product code).
//   A: thread per segment; every 32 outputs the warp transposes a 32x33 tile through shared memory and stores, per
//      segment, 32 consecutive doubles (the flush of the specialised kernels, jit.cpp)
//   B: like A without the shared-memory tile (value made up in the storing lane): isolates the LDS/STS cost
//   C: warp per segment, each warp streams its segment's region front to back
//   D: plain grid-stride streaming store of the whole array (what a memset reaches)
//   E: like A but the loop over the 32 segments is unrolled by 8
// Usage: store_pattern_bench [nseg] [doubles_per_segment] [row_stride] [fma_per_output]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

// row stride (doubles) between the regions of consecutive segments (>= doubles written per segment: a fused piece writes
// a part of every region) and dependent FMAs per output (arithmetic between two stores)
__constant__ int c_stride, c_nf;

__global__ void __launch_bounds__(64) kA(double* __restrict__ out, int nseg, int dps, int nchunk) {
  __shared__ double tile[64 * 33];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int seg = blockIdx.x * 64 + threadIdx.x;
  const int seg0 = seg - lane;
  if (seg0 >= nseg) return;
  const int nvalid = min(32, nseg - seg0);
  double* own = tile + threadIdx.x * 33;
  const double* col = tile + (wib * 32) * 33 + lane;
  double v = seg * 1e-3;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int k = 0; k < 32; k++) {
      for (int q = 0; q < c_nf; q++) v = fma(v, 1.0000001, 0.5);
      own[k] = v;
    }
    __syncwarp();
    double* p = out + (size_t) seg0 * c_stride + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
    for (int s = 0; s < nvalid; s++) {
      const double x = col[s * 33];
      if (ok) p[(size_t) s * c_stride] = x;
    }
    __syncwarp();
  }
}
__global__ void __launch_bounds__(64) kE(double* __restrict__ out, int nseg, int dps, int nchunk) {
  __shared__ double tile[64 * 33];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int seg = blockIdx.x * 64 + threadIdx.x;
  const int seg0 = seg - lane;
  if (seg0 + 32 > nseg) return;
  double* own = tile + threadIdx.x * 33;
  const double* col = tile + (wib * 32) * 33 + lane;
  double v = seg * 1e-3;
  for (int c = 0; c < nchunk; c++) {
#pragma unroll
    for (int k = 0; k < 32; k++) {
      for (int q = 0; q < c_nf; q++) v = fma(v, 1.0000001, 0.5);
      own[k] = v;
    }
    __syncwarp();
    double* p = out + (size_t) seg0 * c_stride + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
#pragma unroll 1
    for (int s = 0; s < 32; s += 8) {
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u++) x[u] = col[(s + u) * 33];
#pragma unroll
      for (int u = 0; u < 8; u++)
        if (ok) p[(size_t) (s + u) * c_stride] = x[u];
    }
    __syncwarp();
  }
}
__global__ void __launch_bounds__(64) kB(double* __restrict__ out, int nseg, int dps, int nchunk) {
  const int lane = threadIdx.x & 31;
  const int seg = blockIdx.x * 64 + threadIdx.x;
  const int seg0 = seg - lane;
  if (seg0 >= nseg) return;
  const int nvalid = min(32, nseg - seg0);
  double v = seg * 1e-3;
  for (int c = 0; c < nchunk; c++) {
    double* p = out + (size_t) seg0 * c_stride + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
    for (int s = 0; s < nvalid; s++) {
      v = fma(v, 1.0000001, 0.5);
      if (ok) p[(size_t) s * c_stride] = v;
    }
  }
}
__global__ void __launch_bounds__(256) kC(double* __restrict__ out, int nseg, int dps) {
  const int lane = threadIdx.x & 31;
  const int w = (blockIdx.x * 256 + threadIdx.x) >> 5;
  if (w >= nseg) return;
  double* p = out + (size_t) w * c_stride;
  double v = w * 1e-3;
  for (int k = lane; k < dps; k += 32) {
    v = fma(v, 1.0000001, 0.5);
    p[k] = v;
  }
}
__global__ void __launch_bounds__(256) kD(double* __restrict__ out, size_t n) {
  const size_t stride = (size_t) gridDim.x * 256;
  for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n; i += stride) out[i] = (double) i;
}

// F<CW>: like A with CW outputs per thread between two walks, so a row receives CW/32 consecutive 256-byte stores per
// walk; dynamic shared memory: blockDim.x rows of CW+1 doubles (+ padding to cap the resident blocks)
template <int CW>
__global__ void __launch_bounds__(32) kF(double* __restrict__ out, int nseg, int dps) {
  extern __shared__ double dyn_tile[];
  const int lane = threadIdx.x & 31;
  const int seg0 = blockIdx.x * 32;
  if (seg0 + 32 > nseg) return;
  double* own = dyn_tile + threadIdx.x * (CW + 1);
  double v = (seg0 + lane) * 1e-3;
  for (int c = 0; c * CW < dps; c++) {
#pragma unroll 8
    for (int k = 0; k < CW; k++) {
      for (int q = 0; q < c_nf; q++) v = fma(v, 1.0000001, 0.5);
      own[k] = v;
    }
    __syncwarp();
    double* p = out + (size_t) seg0 * c_stride + c * CW + lane;
#pragma unroll 1
    for (int s = 0; s < 32; s += 4) {
      double x[4][CW / 32];
#pragma unroll
      for (int u = 0; u < 4; u++)
#pragma unroll
        for (int q = 0; q < CW / 32; q++) x[u][q] = dyn_tile[(s + u) * (CW + 1) + q * 32 + lane];
#pragma unroll
      for (int u = 0; u < 4; u++)
#pragma unroll
        for (int q = 0; q < CW / 32; q++)
          if (c * CW + q * 32 + lane < dps) p[(size_t) (s + u) * c_stride + q * 32] = x[u][q];
    }
    __syncwarp();
  }
}
// G: warp per segment streaming (C) in one-warp blocks, resident blocks capped by dynamic shared memory
__global__ void __launch_bounds__(32) kG(double* __restrict__ out, int nseg, int dps) {
  const int lane = threadIdx.x;
  for (int w = blockIdx.x; w < nseg; w += gridDim.x) {
    double* p = out + (size_t) w * c_stride;
    double v = w * 1e-3;
    for (int k = lane; k < dps; k += 32) {
      for (int q = 0; q < c_nf; q++) v = fma(v, 1.0000001, 0.5);
      p[k] = v;
    }
  }
}

int main(int argc, char** argv) {
  const int nseg = argc > 1 ? atoi(argv[1]) : 200000;
  const int dps = argc > 2 ? atoi(argv[2]) : 910;
  const int nchunk = (dps + 31) / 32;
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
  timeit("A tile-transposed, thread per segment", [&] { kA<<<(nseg + 63) / 64, 64>>>(out, nseg, dps, nchunk); });
  timeit("E same, segment loop unrolled by 8   ", [&] { kE<<<(nseg + 63) / 64, 64>>>(out, nseg, dps, nchunk); });
  // occupancy sweep: dynamic shared memory caps the resident blocks (2 warps each) per SM
  cudaFuncSetAttribute(kA, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kE, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int blocks: {8, 4, 2}) {
    const size_t dyn = (size_t) (220 * 1024) / blocks - 17 * 1024 - 1024;
    char name[96];
    snprintf(name, sizeof(name), "A at %2d warps/SM                       ", 2 * blocks);
    timeit(name, [&] { kA<<<(nseg + 63) / 64, 64, dyn>>>(out, nseg, dps, nchunk); });
    snprintf(name, sizeof(name), "E at %2d warps/SM                       ", 2 * blocks);
    timeit(name, [&] { kE<<<(nseg + 63) / 64, 64, dyn>>>(out, nseg, dps, nchunk); });
  }
  timeit("B no tile, thread per segment        ", [&] { kB<<<(nseg + 63) / 64, 64>>>(out, nseg, dps, nchunk); });
  timeit("C warp per segment, streaming        ", [&] { kC<<<(nseg + 7) / 8, 256>>>(out, nseg, dps); });
  timeit("D grid-stride streaming              ", [&] { kD<<<148 * 8, 256>>>(out, n); });
  // one-warp blocks at a fixed number of resident warps per SM: CW outputs per thread between walks (F), and a warp
  // streaming one segment at a time (G)
  cudaFuncSetAttribute(kF<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kF<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kF<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kG, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int warps: {4, 8, 16}) {
    const size_t dyn = (size_t) (224 * 1024) / warps - 1024;
    char name[96];
    snprintf(name, sizeof(name), "F<32>  one-warp blocks, %2d warps/SM    ", warps);
    timeit(name, [&] { kF<32><<<nseg / 32, 32, dyn>>>(out, nseg, dps); });
    snprintf(name, sizeof(name), "F<64>  one-warp blocks, %2d warps/SM    ", warps);
    if (dyn >= 32 * 65 * 8) timeit(name, [&] { kF<64><<<nseg / 32, 32, dyn>>>(out, nseg, dps); });
    if (dyn >= 32 * 129 * 8) {
      snprintf(name, sizeof(name), "F<128> one-warp blocks, %2d warps/SM    ", warps);
      timeit(name, [&] { kF<128><<<nseg / 32, 32, dyn>>>(out, nseg, dps); });
    }
    snprintf(name, sizeof(name), "G streaming warp, %2d warps/SM          ", warps);
    timeit(name, [&] { kG<<<148 * warps, 32, dyn>>>(out, nseg, dps); });
  }
  return 0;
}
