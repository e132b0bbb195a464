// Micro-benchmark (measurement tool, not product code): tensor memory as lane-private parking space for fp64 values.
//
// The fused KKT kernels keep ~360 doubles alive per thread: 112 in registers, the rest parked in shared memory, which
// caps the residency at three warps per SM.  Blackwell's tensor memory (256 KB per SM, 128 lanes x 512 columns of 32 bit)
// is idle on this path; tcgen05.st / tcgen05.ld with the 32x32b shape give every thread of a warp its own lane, i.e. a
// second lane-private scratch.  This tool checks that it works as one (values come back bit-exact, in any order, with
// and without tcgen05.wait::st in between) and measures what a park / reload costs against shared memory:
//   each thread runs `rounds` rounds of: NV dependent FMAs on NV values, park them all, reload them in reverse order.
// Usage: tmem_park_bench [blocks_per_sm]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int NV = 64;  // doubles parked per thread (128 columns)

template <int MODE>  // 0 registers only, 1 shared memory, 2 tensor memory
__global__ void __launch_bounds__(128) k_park(double* out, int rounds, double a, double b) {
  extern __shared__ double dyn[];
  __shared__ unsigned tm_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  unsigned taddr = 0;
  if (MODE == 2) {
    if (warp == 0) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((unsigned) __cvta_generic_to_shared(&tm_base_s)),
                   "n"(2 * NV));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    taddr = tm_base_s + ((unsigned) (warp * 32) << 16);  // this warp's 32 lanes
  }
  double v[NV];
#pragma unroll
  for (int j = 0; j < NV; j++) v[j] = (blockIdx.x * blockDim.x + tid) * 1e-6 + j;
  double acc = 0.0;
  for (int r = 0; r < rounds; r++) {
#pragma unroll
    for (int j = 0; j < NV; j++) v[j] = fma(v[j], a, b);
    if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < NV; j++) dyn[j * blockDim.x + tid] = v[j];
#pragma unroll
      for (int j = 0; j < NV; j++) v[NV - 1 - j] = dyn[j * blockDim.x + tid] + 1.0;
    } else if (MODE == 2) {
#pragma unroll
      for (int j = 0; j < NV; j++) {
        const unsigned lo = (unsigned) __double2loint(v[j]), hi = (unsigned) __double2hiint(v[j]);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr + 2 * j), "r"(lo), "r"(hi) : "memory");
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < NV; j++) {
        unsigned lo, hi;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(taddr + 2 * j) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        v[NV - 1 - j] = __hiloint2double((int) hi, (int) lo) + 1.0;
      }
    } else {
      double w[NV];
#pragma unroll
      for (int j = 0; j < NV; j++) w[NV - 1 - j] = v[j] + 1.0;
#pragma unroll
      for (int j = 0; j < NV; j++) v[j] = w[j];
    }
  }
#pragma unroll
  for (int j = 0; j < NV; j++) acc += v[j] * (j + 1);
  out[blockIdx.x * blockDim.x + tid] = acc;
  if (MODE == 2) {
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base_s), "n"(2 * NV));
  }
}

int main(int argc, char** argv) {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount, rounds = 200;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int threads: {32, 128}) {
    for (int bps: {1, 2, 3, 4}) {
      if (threads == 32 && bps * 2 * NV > 512) continue;
      if (threads == 128 && bps * 2 * NV > 512) continue;
      const int blocks = nsm * bps * 8;
      const size_t n = (size_t) blocks * threads;
      double *o0, *o1, *o2;
      cudaMalloc(&o0, n * 8);
      cudaMalloc(&o1, n * 8);
      cudaMalloc(&o2, n * 8);
      // residency: blocks per SM capped through dynamic shared memory
      const size_t dyn = (size_t) (220 * 1024) / bps - 2048;
      cudaFuncSetAttribute(k_park<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) dyn);
      cudaFuncSetAttribute(k_park<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) dyn);
      cudaFuncSetAttribute(k_park<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) dyn);
      float ms[3];
      auto run = [&](int mode, double* o) {
        auto launch = [&]() {
          if (mode == 0) k_park<0><<<blocks, threads, dyn>>>(o, rounds, 1.0000001, 0.5);
          if (mode == 1) k_park<1><<<blocks, threads, dyn>>>(o, rounds, 1.0000001, 0.5);
          if (mode == 2) k_park<2><<<blocks, threads, dyn>>>(o, rounds, 1.0000001, 0.5);
        };
        launch();
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms[mode], e0, e1);
      };
      run(0, o0);
      run(1, o1);
      run(2, o2);
      std::vector<double> h0(n), h1(n), h2(n);
      cudaMemcpy(h0.data(), o0, n * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(h1.data(), o1, n * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(h2.data(), o2, n * 8, cudaMemcpyDeviceToHost);
      size_t bad1 = 0, bad2 = 0;
      for (size_t i = 0; i < n; i++) {
        bad1 += h1[i] != h0[i];
        bad2 += h2[i] != h0[i];
      }
      const double parks = (double) n * rounds * NV;
      printf("%3d-thread blocks, %d per SM (%2d warps/SM): registers %.3f ms | shared %.3f ms (%.2f ns/park+reload/SM-warp) | tensor memory "
             "%.3f ms | mismatches shared %zu, tensor %zu of %zu (%s)\n",
             threads, bps, bps * threads / 32, ms[0], ms[1], (ms[1] - ms[0]) * 1e6 / (parks / 32 / nsm), ms[2], bad1, bad2, n,
             cudaGetErrorString(cudaGetLastError()));
      cudaFree(o0);
      cudaFree(o1);
      cudaFree(o2);
    }
  }
  return 0;
}
