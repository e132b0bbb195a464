// Micro-benchmark of the output store patterns of the segment kernels (measurement tool, not product code).
//   A: thread per segment; every 32 outputs the warp transposes a 32x33 tile through shared memory and stores, per
//      segment, 32 consecutive doubles (the flush of the specialised kernels, jit.cpp)
//   B: like A without the shared-memory tile (value made up in the storing lane): isolates the LDS/STS cost
//   C: warp per segment, each warp streams its segment's region front to back
//   D: plain grid-stride streaming store of the whole array (what a memset reaches)
//   E: like A but the loop over the 32 segments is unrolled by 8
// Usage: store_pattern_bench [nseg] [doubles_per_segment]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

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
      v = fma(v, 1.0000001, 0.5);
      own[k] = v;
    }
    __syncwarp();
    double* p = out + (size_t) seg0 * dps + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
    for (int s = 0; s < nvalid; s++) {
      const double x = col[s * 33];
      if (ok) p[(size_t) s * dps] = x;
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
      v = fma(v, 1.0000001, 0.5);
      own[k] = v;
    }
    __syncwarp();
    double* p = out + (size_t) seg0 * dps + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
#pragma unroll 1
    for (int s = 0; s < 32; s += 8) {
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u++) x[u] = col[(s + u) * 33];
#pragma unroll
      for (int u = 0; u < 8; u++)
        if (ok) p[(size_t) (s + u) * dps] = x[u];
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
    double* p = out + (size_t) seg0 * dps + c * 32 + lane;
    const bool ok = c * 32 + lane < dps;
    for (int s = 0; s < nvalid; s++) {
      v = fma(v, 1.0000001, 0.5);
      if (ok) p[(size_t) s * dps] = v;
    }
  }
}
__global__ void __launch_bounds__(256) kC(double* __restrict__ out, int nseg, int dps) {
  const int lane = threadIdx.x & 31;
  const int w = (blockIdx.x * 256 + threadIdx.x) >> 5;
  if (w >= nseg) return;
  double* p = out + (size_t) w * dps;
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

int main(int argc, char** argv) {
  const int nseg = argc > 1 ? atoi(argv[1]) : 200000;
  const int dps = argc > 2 ? atoi(argv[2]) : 910;
  const int nchunk = (dps + 31) / 32;
  const size_t n = (size_t) nseg * dps;
  double* out;
  cudaMalloc(&out, n * 8 + 4096);
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
  printf("nseg %d, %d doubles per segment, %.1f MB\n", nseg, dps, n * 8 / 1e6);
  timeit("A tile-transposed, thread per segment", [&] { kA<<<(nseg + 63) / 64, 64>>>(out, nseg, dps, nchunk); });
  timeit("E same, segment loop unrolled by 8   ", [&] { kE<<<(nseg + 63) / 64, 64>>>(out, nseg, dps, nchunk); });
  // occupancy sweep: dynamic shared memory caps the resident blocks (2 warps each) per SM
  cudaFuncSetAttribute(kA, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(kE, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int blocks: {16, 8, 4, 2}) {
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
  return 0;
}
