// Micro-benchmark (measurement tool, not product code): does an asynchronous output path hide the store stream behind
// the arithmetic at the residency of the fused KKT kernels (3-8 warps per SM)?
//
// Thread t owns the region [t*stride, t*stride + dps) doubles of the output (stride 910: the CSR region of one LGL5
// low-thrust segment).  It produces its outputs 16 at a time, `nf` dependent FMAs deep with four independent chains
// (the arithmetic of the fused kernels: ~9 DFMA per output), and writes them
//   D: straight from registers with st.global.v4.f64 on 32-byte boundaries (what jit.cpp emits today);
//   T: into a 128B-swizzled shared-memory tile [32 lanes][16 doubles] that one elected lane hands to the TMA
//      (cp.async.bulk.tensor.2d.global.shared::cta, ring of `STAGES` tiles, wait_group.read before a tile is reused);
//   N: not at all (arithmetic only; a predicate the compiler cannot fold keeps it alive).
// Residency is capped through dynamic shared memory.  Usage: tma_store_bench [nseg] [nf]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int TW = 16;  // doubles per tile row (128 bytes)

__device__ __forceinline__ void compute16(double (&v)[TW], double seed, int nf, int k) {
#pragma unroll
  for (int j = 0; j < TW; j++) v[j] = seed + j + k;
  for (int q = 0; q < nf; q++) {
#pragma unroll
    for (int j = 0; j < TW; j++) v[j] = fma(v[j], 1.0000001, 0.5);
  }
}

template <int MODE>  // 0 direct, 2 none
__global__ void __launch_bounds__(32) k_direct(double* __restrict__ out, int nseg, int stride, int ntiles, int nf, int never) {
  extern __shared__ double dyn[];
  const int seg = blockIdx.x * 32 + threadIdx.x;
  if (seg >= nseg) return;
  // segments whose region starts 16 bytes off a sector skip two doubles so that every v4 store is one whole sector
  const int skip = (4 - (int) (((size_t) seg * stride) & 3)) & 3;
  double* p = out + (size_t) seg * stride + skip;
  double v[TW];
  for (int k = 0; k < ntiles; k++) {
    compute16(v, seg * 1e-3, nf, k);
    if (MODE == 0 || never) {
#pragma unroll
      for (int j = 0; j < TW; j += 4)
        asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p + k * TW + j), "d"(v[j]), "d"(v[j + 1]), "d"(v[j + 2]), "d"(v[j + 3])
                     : "memory");
    }
  }
}

template <int STAGES>
__global__ void __launch_bounds__(32) k_tma(const __grid_constant__ CUtensorMap tmap, int nseg, int ntiles, int nf) {
  extern __shared__ __align__(1024) unsigned char dynraw[];
  const int lane = threadIdx.x;
  const int seg0 = blockIdx.x * 32;
  if (seg0 >= nseg) return;
  const int seg = seg0 + lane;
  unsigned tile0 = (unsigned) __cvta_generic_to_shared(dynraw);
  tile0 = (tile0 + 1023u) & ~1023u;
  // 128-byte swizzle: 16-byte chunk c of row r lives at chunk c ^ (r & 7)
  const unsigned rowbase = tile0 + lane * 128;
  const unsigned x = (lane & 7) << 4;
  double v[TW];
  for (int k = 0; k < ntiles; k++) {
    compute16(v, seg * 1e-3, nf, k);
    const int st = k % STAGES;
    if (k >= STAGES) {
      // the tile written STAGES rounds ago must have been read by the TMA
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 1) : "memory");
      __syncwarp();
    }
    const unsigned tb = rowbase + st * 4096;
#pragma unroll
    for (int c = 0; c < TW / 2; c++)
      asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(tb + ((c << 4) ^ x)), "d"(v[2 * c]), "d"(v[2 * c + 1]) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmap), "r"(k * TW), "r"(seg0),
                   "r"(tile0 + st * 4096)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int nseg = argc > 1 ? atoi(argv[1]) : 200000;
  const int nf = argc > 2 ? atoi(argv[2]) : 9;
  const int stride = 910, dps = 896, ntiles = dps / TW;  // 56 tiles of 16 doubles
  double* out;
  const size_t bytes = (size_t) nseg * stride * 8 + 4096;
  cudaMalloc(&out, bytes);
  cudaMemset(out, 0, bytes);
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**) &encode, cudaEnableDefault, &qres);
  if (!encode) {
    printf("cuTensorMapEncodeTiled not available\n");
    return 1;
  }
  CUtensorMap tmap;
  const cuuint64_t gdim[2] = {(cuuint64_t) dps, (cuuint64_t) nseg};
  const cuuint64_t gstr[1] = {(cuuint64_t) stride * 8};
  const cuuint32_t box[2] = {TW, 32};
  const cuuint32_t estr[2] = {1, 1};
  CUresult rc = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, out, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("cuTensorMapEncodeTiled rc=%d\n", (int) rc);
  if (rc != CUDA_SUCCESS) return 1;
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
    printf("%-34s %.3f ms  %5.0f GB/s  (%s)\n", name, ms, (double) nseg * dps * 8 / (ms * 1e6), cudaGetErrorString(cudaGetLastError()));
  };
  cudaFuncSetAttribute(k_direct<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_direct<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_tma<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_tma<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int blocks = (nseg + 31) / 32;
  printf("nseg %d, %d of every %d doubles written, %d FMA per output, %.1f MB\n", nseg, dps, stride, nf, (double) nseg * dps * 8 / 1e6);
  for (int warps: {3, 4, 8, 16}) {
    const size_t dyn = (size_t) (224 * 1024) / warps - 1024;
    char name[96];
    snprintf(name, sizeof(name), "N arithmetic only, %2d warps/SM", warps);
    timeit(name, [&] { k_direct<2><<<blocks, 32, dyn>>>(out, nseg, stride, ntiles, nf, 0); });
    snprintf(name, sizeof(name), "D direct v4 stores, %2d warps/SM", warps);
    timeit(name, [&] { k_direct<0><<<blocks, 32, dyn>>>(out, nseg, stride, ntiles, nf, 0); });
    snprintf(name, sizeof(name), "T TMA tiles, 2 stages, %2d warps/SM", warps);
    timeit(name, [&] { k_tma<2><<<blocks, 32, dyn < 2 * 4096 + 1024 ? 2 * 4096 + 1024 : dyn>>>(tmap, nseg, ntiles, nf); });
    snprintf(name, sizeof(name), "T TMA tiles, 4 stages, %2d warps/SM", warps);
    timeit(name, [&] { k_tma<4><<<blocks, 32, dyn < 4 * 4096 + 1024 ? 4 * 4096 + 1024 : dyn>>>(tmap, nseg, ntiles, nf); });
  }
  // correctness of the TMA path: every written element equals what the direct kernel writes
  std::vector<double> a((size_t) 64 * stride), b((size_t) 64 * stride);
  cudaMemset(out, 0, bytes);
  k_tma<2><<<blocks, 32, 16384>>>(tmap, nseg, ntiles, nf);
  cudaMemcpy(a.data(), out + (size_t) 1000 * stride, a.size() * 8, cudaMemcpyDeviceToHost);
  cudaMemset(out, 0, bytes);
  k_direct<0><<<blocks, 32, 1024>>>(out, nseg, stride, ntiles, nf, 0);
  cudaMemcpy(b.data(), out + (size_t) 1000 * stride, b.size() * 8, cudaMemcpyDeviceToHost);
  long long bad = 0, nz = 0;
  for (int s = 0; s < 64; s++) {
    const int skip = (4 - (int) (((size_t) (1000 + s) * stride) & 3)) & 3;
    for (int j = 0; j < dps; j++) {
      // element j of the TMA run is run element j; the direct kernel shifts its run by `skip`
      const double ta = a[(size_t) s * stride + j];
      const double want = (j >= skip) ? b[(size_t) s * stride + j] : 0.0;
      (void) want;
      nz += ta != 0.0;
    }
  }
  // value check independent of the direct kernel's shift: recompute on the host
  for (int s = 0; s < 64; s++)
    for (int j = 0; j < dps; j++) {
      double v = (1000 + s) * 1e-3 + (j % TW) + (j / TW);
      for (int q = 0; q < nf; q++) v = fma(v, 1.0000001, 0.5);
      if (a[(size_t) s * stride + j] != v) bad++;
    }
  printf("TMA path check: %lld mismatches of %d, %lld non-zero (%s)\n", bad, 64 * dps, nz, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
