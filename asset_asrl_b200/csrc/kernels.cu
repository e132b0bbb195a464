// sm_100a fp64 kernels for the KKT evaluation hot path.
//
// Replaces the per-application loop of the reference
// (/root/reference/src/VectorFunctions/DenseFunctionBase.h:1276-1391: gather -> compute -> KKTFillAll)
// and the thread-pool fan-out of NonLinearProgram::evalKKT (NonLinearProgram.cpp:473-537).
//
// k_eval_prog: interpreter of the scalar device program (expr.hpp: Program).  One thread evaluates one function
// application (mesh segment); the 32 lanes of a warp run the same instruction on 32 segments, so instruction fetch
// and decode are warp-uniform and amortised, and there is no divergence.  Every live value sits in a lane-private
// 8-byte shared-memory slot laid out [slot][thread] (bank-conflict free); constants are shared by the block.
// The program already contains the symbolically differentiated, sparsity-pruned arithmetic for values, Jacobian,
// adjoint gradient and lambda-weighted Hessian; almost all of it is fused multiply-adds issued in bundles of four
// independent operations (12 shared loads in flight, 4 DFMA, 4 shared stores).
#include <cuda_runtime.h>
#include <math.h>
#include <algorithm>
#include <stdint.h>

#include "device_types.h"
#include "expr.hpp"

namespace ab {

__device__ __forceinline__ double sgn(double x) { return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : 0.0); }

// operand word -> shared-memory byte address (relative to the dynamic smem base)
__device__ __forceinline__ unsigned op_addr(unsigned w, unsigned tid8) { return (w & ~7u) + ((0u - (w & 1u)) & tid8); }

__device__ __forceinline__ double lds(const unsigned char* smem, unsigned addr) {
  return *reinterpret_cast<const double*>(smem + addr);
}
// load an operand and apply its negate flag (bit 1) by flipping the sign bit
__device__ __forceinline__ double ld_opnd(const unsigned char* smem, unsigned w, unsigned tid8) {
  const double v = lds(smem, op_addr(w, tid8));
  const long long bits = __double_as_longlong(v) ^ ((long long) (w & 2u) << 62);
  return __longlong_as_double(bits);
}

__device__ __forceinline__ void put(const EvalIO& io, const int* __restrict__ dest, long long t, double v) {
  const int d = __ldg(dest + t);
  if (d >= 0)
    io.kkt[d] = v;
  else
    io.stage[~d] = v;
}

// interval of an axis stored as [n, evenly spaced, coordinates...] that holds t (InterpTable1D::get_telem,
// /root/reference/src/VectorFunctions/CommonFunctions/InterpTable1D.h:180-197); same arithmetic as the specialised kernels
__device__ __forceinline__ double table_interval_dev(const double* __restrict__ axis, double t) {
  const int n = (int) __ldg(axis);
  const double* __restrict__ ts = axis + 2;
  int e;
  if (__ldg(axis + 1) != 0.0) {
    const double t0 = __ldg(ts);
    e = __double2int_rz(__ddiv_rn(__dsub_rn(t, t0), __dsub_rn(__ldg(ts + 1), t0)));
  } else {
    // number of coordinates <= t, by descending powers of two
    int pos = 0;
    for (int step = 1 << (31 - __clz(n)); step > 0; step >>= 1) {
      const int cand = pos + step;
      if (cand <= n && __ldg(ts + cand - 1) <= t) pos = cand;
    }
    e = pos - 1;
  }
  return (double) max(min(e, n - 2), 0);
}

__device__ __forceinline__ double special_op(unsigned op, double a, double b, double c, double imm, const double* __restrict__ tab) {
  switch (op) {
    case OP_TABIDX: return table_interval_dev(tab + (long long) imm, a);
    case OP_TABLOAD: return __ldg(tab + (long long) imm + __double2int_rz(a));
    case OP_DIV: return a / b;
    case OP_INV: return 1.0 / a;
    case OP_SQRT: return sqrt(a);
    case OP_EXP: return exp(a);
    case OP_LOG: return log(a);
    case OP_SIN: return sin(a);
    case OP_COS: return cos(a);
    case OP_TAN: return tan(a);
    case OP_ASIN: return asin(a);
    case OP_ACOS: return acos(a);
    case OP_ATAN: return atan(a);
    case OP_SINH: return sinh(a);
    case OP_COSH: return cosh(a);
    case OP_TANH: return tanh(a);
    case OP_ASINH: return asinh(a);
    case OP_ACOSH: return acosh(a);
    case OP_ATANH: return atanh(a);
    case OP_ABS: return fabs(a);
    case OP_SIGN: return sgn(a);
    case OP_POWC: return pow(a, imm);
    case OP_ATAN2: return atan2(a, b);
    case OP_LT: return (a < b) ? 1.0 : 0.0;
    case OP_LE: return (a <= b) ? 1.0 : 0.0;
    case OP_SELECT: return (a != 0.0) ? b : c;
    default: return 0.0;
  }
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const unsigned d = (unsigned) __cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// shared memory: [constants][2 x program chunk][lane-private slots]; operand offsets are linked against this layout
__global__ void __launch_bounds__(256) k_eval_prog(const FnDev f, const EvalIO io) {
  extern __shared__ __align__(16) unsigned char smem[];
  const ProgDev& P = f.prog;
  {
    double* sc = reinterpret_cast<double*>(smem);
    for (int i = threadIdx.x; i < P.nconst; i += blockDim.x) sc[i] = P.consts[i];
  }
  const unsigned const_bytes = ((unsigned) P.nconst * 8u + 15u) & ~15u;
  uint4* const cbuf = reinterpret_cast<uint4*>(smem + const_bytes);  // 2 x PROG_CHUNK_BUNDLES x 4 uint4
  constexpr int CH = (int) PROG_CHUNK_BUNDLES;
  const int nb = P.nbundles;
  const int nchunks = (nb + CH - 1) / CH;
  const unsigned tid8 = threadIdx.x * 8u;
  const int ir = f.ir, orows = f.orows;
  const long long napp = (long long) f.app_end - f.app_begin;
  const long long ntiles = (napp + blockDim.x - 1) / blockDim.x;
  const uint4* __restrict__ code = P.code;
  auto fetch_chunk = [&](int c) {
    const int n16 = min(CH, nb - c * CH) * 4;
    const uint4* src = code + (size_t) c * CH * 4;
    uint4* dst = cbuf + (c & 1) * CH * 4;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) cp_async16(dst + i, src + i);
  };

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    long long app = f.app_begin + tile * blockDim.x + threadIdx.x;
    const bool active = app < f.app_end;
    if (!active) app = f.app_end - 1;
    const bool aff = app >= f.aff_begin && app < f.aff_end;
    const int rel = (int) (app - f.aff_begin);
    const int* __restrict__ vidx = f.vindex + app * ir;
    const int* __restrict__ cidx = f.cindex + app * orows;
    const double* __restrict__ appc = f.appconst + app * f.nappconst;
    const int* __restrict__ kdest = f.kkt_dest + app * f.E;

    __syncthreads();  // every thread is done with the previous tile's last chunk (and the constants are visible)
    fetch_chunk(0);
    for (int c = 0; c < nchunks; c++) {
      cp_async_wait_all();
      __syncthreads();  // chunk c is visible; every thread has finished chunk c-1, whose buffer is refilled next
      if (c + 1 < nchunks) fetch_chunk(c + 1);
      const uint4* __restrict__ cb = cbuf + (c & 1) * CH * 4;
      const int nbc = min(CH, nb - c * CH);
      // software pipeline: the next bundle is read from shared memory while the current one executes
      uint4 n0 = cb[0], n1 = cb[1], n2 = cb[2], n3 = cb[3];
      for (int pc = 1; pc <= nbc; pc++) {
        const uint4 i0 = n0, i1 = n1, i2 = n2, i3 = n3;
        const unsigned kind = i0.x >> 28;
        if (kind == PK_END) break;
        if (pc < nbc) {
          n0 = cb[4 * pc + 0];
          n1 = cb[4 * pc + 1];
          n2 = cb[4 * pc + 2];
          n3 = cb[4 * pc + 3];
        }
        const unsigned cnt = (i0.x >> 24) & 15u;
        if (kind == PK_FMA) {
          // unused instructions of a short bundle are encoded as scratch = 0 * 0 + 0 (see link_program), so all four run
          const double a0 = ld_opnd(smem, i0.y, tid8), b0 = lds(smem, op_addr(i0.z, tid8)), c0 = ld_opnd(smem, i0.w, tid8);
          const double a1 = ld_opnd(smem, i1.y, tid8), b1 = lds(smem, op_addr(i1.z, tid8)), c1 = ld_opnd(smem, i1.w, tid8);
          const double a2 = ld_opnd(smem, i2.y, tid8), b2 = lds(smem, op_addr(i2.z, tid8)), c2 = ld_opnd(smem, i2.w, tid8);
          const double a3 = ld_opnd(smem, i3.y, tid8), b3 = lds(smem, op_addr(i3.z, tid8)), c3 = ld_opnd(smem, i3.w, tid8);
          const double r0 = fma(a0, b0, c0), r1 = fma(a1, b1, c1), r2 = fma(a2, b2, c2), r3 = fma(a3, b3, c3);
          *reinterpret_cast<double*>(smem + op_addr(i0.x & 0x00FFFFFFu, tid8)) = r0;
          *reinterpret_cast<double*>(smem + op_addr(i1.x, tid8)) = r1;
          *reinterpret_cast<double*>(smem + op_addr(i2.x, tid8)) = r2;
          *reinterpret_cast<double*>(smem + op_addr(i3.x, tid8)) = r3;
        } else if (kind == PK_OUT) {
          const uint4 ins[4] = {i0, i1, i2, i3};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (q < cnt) {
              const double v = ld_opnd(smem, ins[q].y, tid8);
              const unsigned ok = ins[q].z & 15u, idx = ins[q].z >> 4;
              const int base = (int) ins[q].w, stride = ((int) (ins[q].x << 8)) >> 8;
              if (active) {
                if (ok == PO_G) {
                  io.stage[f.g_stage + app * ir + idx] = v;
                } else if (ok == PO_F && f.role == ROLE_OBJ) {
                  io.stage[f.v_stage + app] = v * io.obj_scale;
                } else {
                  int d;
                  if (aff)
                    d = base + rel * stride;
                  else
                    d = (ok == PO_F) ? __ldg(f.f_dest + app * orows + idx) : __ldg(kdest + idx);
                  double* tgt = (ok == PO_F) ? io.fx : io.kkt;
                  if (d >= 0)
                    tgt[d] = v;
                  else
                    io.stage[~d] = v;
                }
              }
            }
          }
        } else if (kind == PK_LOAD) {
          const uint4 ins[4] = {i0, i1, i2, i3};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (q < cnt) {
              const unsigned idx = ins[q].y & 0x0FFFFFFFu, ty = ins[q].y >> 28;
              const int gi = (int) ins[q].z + rel * (int) ins[q].w;
              double v;
              if (ty == PL_X)
                v = __ldg(io.X + (aff ? gi : __ldg(vidx + idx)));
              else if (ty == PL_LAMBDA)
                v = (f.role == ROLE_OBJ) ? io.obj_scale : __ldg(io.L + (aff ? gi : __ldg(cidx + idx)));
              else
                v = __ldg(appc + idx);
              *reinterpret_cast<double*>(smem + op_addr(ins[q].x & 0x00FFFFFFu, tid8)) = v;
            }
          }
        } else {  // PK_SPECIAL: one operation
          const unsigned op = i0.w & 0xFFu;
          const double a = ld_opnd(smem, i0.y, tid8);
          const double b = ld_opnd(smem, i0.z, tid8);
          const double c3 = ld_opnd(smem, i1.y, tid8);
          const double imm = lds(smem, (i0.w >> 8) * 8u);
          *reinterpret_cast<double*>(smem + op_addr(i0.x & 0x00FFFFFFu, tid8)) = special_op(op, a, b, c3, imm, P.tables);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Deterministic segmented summation (replaces the mutex-guarded `mpt[lpt[k]] += v` scatter and the
// serial RHSFillOP, NonLinearProgram.h:401-407).  Sources are summed in ascending source order, the
// order RHSFillOP uses.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gather_small(const GatherDev g, const double* __restrict__ stage,
                                                      double* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= g.ntargets) return;
  const long long b = g.ptr[t], e = g.ptr[t + 1];
  double s = 0.0;
  for (long long k = b; k < e; k++) s += stage[g.src[k]];
  out[g.target[t]] = s;
}

// High fan-in targets (objective value, parameter rows, accumulated constraints): two stages, both with a fixed
// partition and a fixed summation tree, so the result does not depend on scheduling.  Stage 1: block (part, target)
// sums sources [part*GATHER_PART, (part+1)*GATHER_PART) of its target; stage 2: one block per target sums the partials.
__device__ __forceinline__ double block_tree_sum(double s, double* part) {
  part[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) part[threadIdx.x] += part[threadIdx.x + w];
    __syncthreads();
  }
  return part[0];
}
__global__ void __launch_bounds__(256) k_gather_large_parts(const GatherDev g, const double* __restrict__ stage,
                                                            double* __restrict__ partials) {
  __shared__ double part[256];
  const int t = blockIdx.y;
  const long long b = g.ptr[t], e = g.ptr[t + 1];
  const long long lo0 = b + (long long) blockIdx.x * GATHER_PART;
  if (lo0 >= e) return;
  const long long hi0 = (lo0 + GATHER_PART < e) ? lo0 + GATHER_PART : e;
  constexpr int PER = GATHER_PART / 256;
  const long long lo = lo0 + (long long) threadIdx.x * PER;
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < PER; k++)
    if (lo + k < hi0) s += stage[g.src[lo + k]];
  s = block_tree_sum(s, part);
  if (threadIdx.x == 0) partials[(size_t) t * g.maxparts + blockIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_gather_large_final(const GatherDev g, const double* __restrict__ partials,
                                                            double* __restrict__ out) {
  __shared__ double part[256];
  const int t = blockIdx.x;
  const long long n = g.ptr[t + 1] - g.ptr[t];
  const int np = (int) ((n + GATHER_PART - 1) / GATHER_PART);
  const int chunk = (np + 255) / 256;
  double s = 0.0;
  for (int k = 0; k < chunk; k++) {
    const int q = threadIdx.x * chunk + k;
    if (q < np) s += partials[(size_t) t * g.maxparts + q];
  }
  s = block_tree_sum(s, part);
  if (threadIdx.x == 0) out[g.target[t]] = s;
}

// ---------------------------------------------------------------------------------------------
// host launchers (called from plan.cu)
// ---------------------------------------------------------------------------------------------
// per device (the C ABI lets one process hold plans on several devices): SM count, and the dynamic shared memory
// k_eval_prog has been allowed so far (cudaFuncSetAttribute is a per-device setting)
constexpr int MAX_DEVICES = 64;
static int g_num_sms[MAX_DEVICES] = {0};
static unsigned g_smem_configured[MAX_DEVICES] = {0};
static int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return (dev >= 0 && dev < MAX_DEVICES) ? dev : 0;
}
static int num_sms() {
  const int dev = current_device();
  if (!g_num_sms[dev]) {
    cudaDeviceGetAttribute(&g_num_sms[dev], cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms[dev] <= 0) g_num_sms[dev] = 148;
  }
  return g_num_sms[dev];
}

cudaError_t launch_eval(const FnDev& f, cudaStream_t st, const EvalIO& io) {
  const long long napp = (long long) f.app_end - f.app_begin;
  if (napp <= 0) return cudaSuccess;
  const ProgDev& P = f.prog;
  unsigned& configured = g_smem_configured[current_device()];
  if (P.smem_bytes > configured) {
    cudaError_t e = cudaFuncSetAttribute(k_eval_prog, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) P.smem_bytes);
    if (e != cudaSuccess) return e;
    configured = P.smem_bytes;
  }
  const long long ntiles = (napp + P.nthreads - 1) / P.nthreads;
  // resident blocks per SM are bounded by shared memory (227 KB) and by 2048 threads; persistent grid-stride beyond that
  long long per_sm = std::min<long long>((227 * 1024) / std::max(P.smem_bytes + 1024u, 1u), 2048 / P.nthreads);
  per_sm = std::max<long long>(per_sm, 1);
  const long long blocks = std::min<long long>(ntiles, (long long) num_sms() * per_sm);
  k_eval_prog<<<(unsigned) blocks, P.nthreads, P.smem_bytes, st>>>(f, io);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_cut_pack(const double* __restrict__ arr, const int* __restrict__ target,
                                                  const unsigned char* __restrict__ flag, long long n, int active,
                                                  double* __restrict__ buf) {
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) buf[i] = (active && (flag[i] & 1)) ? arr[target[i]] : 0.0;
}
__global__ void __launch_bounds__(256) k_cut_unpack(double* __restrict__ arr, const int* __restrict__ target,
                                                    const unsigned char* __restrict__ flag, long long n,
                                                    const double* __restrict__ buf) {
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i] & 2)
    arr[target[i]] = buf[i];
  else if (flag[i] & 1)
    arr[target[i]] = 0.0;
}
__global__ void __launch_bounds__(256) k_clear(double* __restrict__ arr, const int* __restrict__ idx, long long n) {
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) arr[idx[i]] = 0.0;
}
cudaError_t launch_clear(double* arr, const int* idx, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  k_clear<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(arr, idx, n);
  return cudaGetLastError();
}
// all output arrays' cut entries in one launch (the cut set is a handful of doubles: one launch instead of one per array)
__global__ void __launch_bounds__(256) k_cut_all(CutAll c, int pack) {
  const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.off[6]) return;
  int a = 0;
  while (i >= c.off[a + 1]) a++;
  const long long q = i - c.off[a];
  const unsigned char fl = c.flag[a][q];
  double* arr = c.arr[a];
  const int t = c.target[a][q];
  if (pack) {
    c.buf[i] = (c.active[a] && (fl & 1)) ? arr[t] : 0.0;
  } else if (c.active[a]) {
    if (fl & 2) arr[t] = c.buf[i];
    else if (fl & 1) arr[t] = 0.0;
  }
}
cudaError_t launch_cut_all(const CutAll& c, bool pack, cudaStream_t st) {
  if (c.off[6] <= 0) return cudaSuccess;
  k_cut_all<<<(unsigned) ((c.off[6] + 255) / 256), 256, 0, st>>>(c, pack ? 1 : 0);
  return cudaGetLastError();
}

cudaError_t launch_cut(bool pack, bool active, double* arr, const int* target, const unsigned char* flag, long long n,
                       double* buf, cudaStream_t st) {
  const unsigned blocks = (unsigned) ((n + 255) / 256);
  if (pack)
    k_cut_pack<<<blocks, 256, 0, st>>>(arr, target, flag, n, active ? 1 : 0, buf);
  else
    k_cut_unpack<<<blocks, 256, 0, st>>>(arr, target, flag, n, buf);
  return cudaGetLastError();
}

cudaError_t launch_gather(const GatherDev& small, const GatherDev& large, const double* stage, double* out,
                          double* scratch, cudaStream_t st) {
  if (small.ntargets > 0) {
    k_gather_small<<<(small.ntargets + 255) / 256, 256, 0, st>>>(small, stage, out);
  }
  if (large.ntargets > 0) {
    k_gather_large_parts<<<dim3((unsigned) large.maxparts, (unsigned) large.ntargets), 256, 0, st>>>(large, stage, scratch);
    k_gather_large_final<<<large.ntargets, 256, 0, st>>>(large, scratch, out);
  }
  return cudaGetLastError();
}

}  // namespace ab
