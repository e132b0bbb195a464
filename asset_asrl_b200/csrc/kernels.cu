// sm_100a fp64 kernels for the KKT evaluation hot path.
//
// Replaces the per-application loop of the reference
// (/root/reference/src/VectorFunctions/DenseFunctionBase.h:1276-1391: gather -> compute -> KKTFillAll)
// and the thread-pool fan-out of NonLinearProgram::evalKKT (NonLinearProgram.cpp:473-537).
//
// k_eval_dual: the generic tape interpreter.  Lanes are (application, tangent direction) pairs: a function
// with IR inputs packs floor(32/IR) applications into one warp; lane `dir` carries d/dx_dir of every
// slot.  The primal section yields f and the Jacobian column, the adjoint section (symbolic reverse
// sweep emitted on the host) yields the column of the lambda-weighted Hessian.  Slot storage is
// lane-private shared memory [slot][lane] so no warp synchronisation is needed inside the tape loop.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "device_types.h"

namespace ab {

struct __align__(16) Dual {
  double v, d;
};

__device__ __forceinline__ double sgn(double x) { return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : 0.0); }

// One interpreted instruction on a dual number.  Returns false for OUT ops (nothing to store).
template<bool DUAL>
__device__ __forceinline__ Dual exec_op(const Instr& ins, const Dual* __restrict__ slots, int lane, int dir,
                                        const double* __restrict__ X, const int* __restrict__ vidx,
                                        const double* __restrict__ L, const int* __restrict__ cidx, int role,
                                        double obj_scale, const double* __restrict__ appc) {
  Dual r;
  r.d = 0.0;
  const int op = ins.op;
  if (op <= OP_APPCONST) {
    switch (op) {
      case OP_CONST: r.v = ins.imm; break;
      case OP_INPUT:
        r.v = __ldg(X + vidx[ins.a]);
        if (DUAL) r.d = (ins.a == dir) ? 1.0 : 0.0;
        break;
      case OP_LAMBDA: r.v = (role == ROLE_OBJ) ? obj_scale : __ldg(L + cidx[ins.a]); break;
      default: r.v = appc[ins.a]; break;
    }
    return r;
  }
  const Dual a = slots[(int) ins.a * 32 + lane];
  switch (op) {
    case OP_ADD: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = a.v + b.v;
      if (DUAL) r.d = a.d + b.d;
    } break;
    case OP_SUB: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = a.v - b.v;
      if (DUAL) r.d = a.d - b.d;
    } break;
    case OP_MUL: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = a.v * b.v;
      if (DUAL) r.d = fma(a.d, b.v, a.v * b.d);
    } break;
    case OP_DIV: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = a.v / b.v;
      if (DUAL) r.d = (a.d - r.v * b.d) / b.v;
    } break;
    case OP_NEG:
      r.v = -a.v;
      if (DUAL) r.d = -a.d;
      break;
    case OP_ADDC:
      r.v = a.v + ins.imm;
      if (DUAL) r.d = a.d;
      break;
    case OP_MULC:
      r.v = a.v * ins.imm;
      if (DUAL) r.d = a.d * ins.imm;
      break;
    case OP_SQUARE:
      r.v = a.v * a.v;
      if (DUAL) r.d = 2.0 * a.v * a.d;
      break;
    case OP_INV:
      r.v = 1.0 / a.v;
      if (DUAL) r.d = -a.d * r.v * r.v;
      break;
    case OP_SQRT:
      r.v = sqrt(a.v);
      if (DUAL) r.d = 0.5 * a.d / r.v;
      break;
    case OP_EXP:
      r.v = exp(a.v);
      if (DUAL) r.d = r.v * a.d;
      break;
    case OP_LOG:
      r.v = log(a.v);
      if (DUAL) r.d = a.d / a.v;
      break;
    case OP_SIN: {
      double s, c;
      sincos(a.v, &s, &c);
      r.v = s;
      if (DUAL) r.d = c * a.d;
    } break;
    case OP_COS: {
      double s, c;
      sincos(a.v, &s, &c);
      r.v = c;
      if (DUAL) r.d = -s * a.d;
    } break;
    case OP_TAN:
      r.v = tan(a.v);
      if (DUAL) r.d = a.d * fma(r.v, r.v, 1.0);
      break;
    case OP_ASIN:
      r.v = asin(a.v);
      if (DUAL) r.d = a.d / sqrt(1.0 - a.v * a.v);
      break;
    case OP_ACOS:
      r.v = acos(a.v);
      if (DUAL) r.d = -a.d / sqrt(1.0 - a.v * a.v);
      break;
    case OP_ATAN:
      r.v = atan(a.v);
      if (DUAL) r.d = a.d / fma(a.v, a.v, 1.0);
      break;
    case OP_SINH:
      r.v = sinh(a.v);
      if (DUAL) r.d = cosh(a.v) * a.d;
      break;
    case OP_COSH:
      r.v = cosh(a.v);
      if (DUAL) r.d = sinh(a.v) * a.d;
      break;
    case OP_TANH:
      r.v = tanh(a.v);
      if (DUAL) r.d = a.d * (1.0 - r.v * r.v);
      break;
    case OP_ASINH:
      r.v = asinh(a.v);
      if (DUAL) r.d = a.d / sqrt(fma(a.v, a.v, 1.0));
      break;
    case OP_ACOSH:
      r.v = acosh(a.v);
      if (DUAL) r.d = a.d / sqrt(fma(a.v, a.v, -1.0));
      break;
    case OP_ATANH:
      r.v = atanh(a.v);
      if (DUAL) r.d = a.d / (1.0 - a.v * a.v);
      break;
    case OP_ABS:
      r.v = fabs(a.v);
      if (DUAL) r.d = sgn(a.v) * a.d;
      break;
    case OP_SIGN: r.v = sgn(a.v); break;
    case OP_POWC:
      r.v = pow(a.v, ins.imm);
      if (DUAL) r.d = ins.imm * pow(a.v, ins.imm - 1.0) * a.d;
      break;
    case OP_ATAN2: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = atan2(a.v, b.v);
      if (DUAL) r.d = (b.v * a.d - a.v * b.d) / fma(a.v, a.v, b.v * b.v);
    } break;
    case OP_LT: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = (a.v < b.v) ? 1.0 : 0.0;
    } break;
    case OP_LE: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      r.v = (a.v <= b.v) ? 1.0 : 0.0;
    } break;
    case OP_SELECT: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      const Dual c = slots[(int) ins.c * 32 + lane];
      r = (a.v != 0.0) ? b : c;
    } break;
    case OP_FMA: {
      const Dual b = slots[(int) ins.b * 32 + lane];
      const Dual c = slots[(int) ins.c * 32 + lane];
      r.v = fma(a.v, b.v, c.v);
      if (DUAL) r.d = fma(a.d, b.v, fma(a.v, b.d, c.d));
    } break;
    default: r.v = 0.0; break;
  }
  return r;
}

__device__ __forceinline__ void put(const EvalIO& io, const int* __restrict__ dest, long long t, double v) {
  const int d = __ldg(dest + t);
  if (d >= 0)
    io.kkt[d] = v;
  else
    io.stage[~d] = v;
}

// Triplet position of the start of input column i inside one application's block, following the
// emission order of DenseFunctionBase::getKKTSpace (DenseFunctionBase.h:1097-1129):
// for each i: [lower-triangular Hessian rows j>=i] then [Jacobian rows 0..OR).
__device__ __forceinline__ int col_base(int i, int ir, int orows, bool hess, bool jac) {
  int b = 0;
  if (hess) b += i * ir - (i * (i - 1)) / 2;
  if (jac) b += i * orows;
  return b;
}

template<int MODE>
__global__ void __launch_bounds__(128) k_eval_dual(const FnDev f, const EvalIO io) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr bool WANT_H = (MODE == FM_ALL);
  const bool want_h = WANT_H && !f.linear;
  const int nins = want_h ? f.n_total : f.n_primal;
  const int nslots = want_h ? f.slots_total : f.slots_primal;
  Instr* stape = reinterpret_cast<Instr*>(smem_raw);
  for (int i = threadIdx.x; i < nins; i += blockDim.x) stape[i] = f.tape[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  Dual* slots = reinterpret_cast<Dual*>(smem_raw + (size_t) nins * sizeof(Instr)) + (size_t) wib * nslots * 32;

  const int ir = f.ir, orows = f.orows;
  const int apw = (ir <= 32) ? (32 / ir) : 1;           // applications per warp
  const int npass = (ir <= 32) ? 1 : (ir + 31) / 32;    // tangent-direction passes
  const int napp = f.app_end - f.app_begin;
  const int nitems = (napp + apw - 1) / apw;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool kkt_jac = (MODE >= FM_JAC) && (f.role != ROLE_OBJ);
  const bool want_g = (MODE == FM_GRAD || MODE == FM_JAC_GRAD || MODE == FM_ALL);

  for (int item = gwarp; item < nitems; item += nwarps) {
    for (int pass = 0; pass < npass; pass++) {
      int grp, dir;
      if (ir <= 32) {
        grp = lane / ir;
        dir = lane - grp * ir;
      } else {
        grp = 0;
        dir = pass * 32 + lane;
      }
      int app = f.app_begin + item * apw + grp;
      const bool active = (grp < apw) && (app < f.app_end) && (dir < ir);
      if (app >= f.app_end) app = f.app_end - 1;
      const int* vidx = f.vindex + (size_t) app * ir;
      const int* cidx = f.cindex + (size_t) app * orows;
      const double* appc = f.appconst + (size_t) app * f.nappconst;
      // triplet layout is fixed by the function type, not by what this mode computes
      const bool lay_h = !f.linear, lay_j = (f.role != ROLE_OBJ);
      const long long tbase = (long long) app * f.E + col_base(dir, ir, orows, lay_h, lay_j);
      const int hcount = lay_h ? (ir - dir) : 0;
      double gacc = 0.0;

      for (int pc = 0; pc < nins; ++pc) {
        const Instr ins = stape[pc];
        if (ins.op == OP_OUT_F) {
          const Dual a = slots[(int) ins.a * 32 + lane];
          const int k = ins.dst;
          if (active) {
            if (dir == 0) {
              if (f.role == ROLE_OBJ) {
                io.stage[f.v_stage + app] = a.v * io.obj_scale;
              } else {
                const int fd = __ldg(f.f_dest + (size_t) app * orows + k);
                if (fd >= 0)
                  io.fx[fd] = a.v;
                else
                  io.stage[~fd] = a.v;
              }
            }
            if (want_g) {
              const double lam = (f.role == ROLE_OBJ) ? io.obj_scale : __ldg(io.L + cidx[k]);
              gacc = fma(lam, a.d, gacc);
            }
            if (kkt_jac) put(io, f.kkt_dest, tbase + hcount + k, a.d);
          }
        } else if (ins.op == OP_OUT_G) {
          const Dual a = slots[(int) ins.a * 32 + lane];
          const int j = ins.dst;
          if (active && j >= dir) put(io, f.kkt_dest, tbase + (j - dir), a.d);
        } else {
          slots[(int) ins.dst * 32 + lane] =
              exec_op<true>(ins, slots, lane, dir, io.X, vidx, io.L, cidx, f.role, io.obj_scale, appc);
        }
      }
      if (want_g && active) io.stage[f.g_stage + (long long) app * ir + dir] = gacc;
    }
  }
}

// Values only: one lane per application, no tangents.
__global__ void __launch_bounds__(128) k_eval_values(const FnDev f, const EvalIO io) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nins = f.n_primal;
  const int nslots = f.slots_primal;
  Instr* stape = reinterpret_cast<Instr*>(smem_raw);
  for (int i = threadIdx.x; i < nins; i += blockDim.x) stape[i] = f.tape[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  Dual* slots = reinterpret_cast<Dual*>(smem_raw + (size_t) nins * sizeof(Instr)) + (size_t) wib * nslots * 32;
  const int napp = f.app_end - f.app_begin;
  const int stride = gridDim.x * blockDim.x;
  const int nround = (napp + stride - 1) / stride;
  for (int r = 0; r < nround; r++) {
    int app = f.app_begin + r * stride + blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = app < f.app_end;
    if (!active) app = f.app_end - 1;
    const int* vidx = f.vindex + (size_t) app * f.ir;
    const int* cidx = f.cindex + (size_t) app * f.orows;
    const double* appc = f.appconst + (size_t) app * f.nappconst;
    for (int pc = 0; pc < nins; ++pc) {
      const Instr ins = stape[pc];
      if (ins.op == OP_OUT_F) {
        const double v = slots[(int) ins.a * 32 + lane].v;
        if (active) {
          if (f.role == ROLE_OBJ) {
            io.stage[f.v_stage + app] = v * io.obj_scale;
          } else {
            const int fd = __ldg(f.f_dest + (size_t) app * f.orows + ins.dst);
            if (fd >= 0)
              io.fx[fd] = v;
            else
              io.stage[~fd] = v;
          }
        }
      } else {
        slots[(int) ins.dst * 32 + lane] =
            exec_op<false>(ins, slots, lane, -1, io.X, vidx, io.L, cidx, f.role, io.obj_scale, appc);
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

// High fan-in targets (objective value, parameter rows): one block per target, fixed partition and
// fixed tree so the result does not depend on scheduling.
__global__ void __launch_bounds__(256) k_gather_large(const GatherDev g, const double* __restrict__ stage,
                                                      double* __restrict__ out) {
  __shared__ double part[256];
  const int t = blockIdx.x;
  const long long b = g.ptr[t], e = g.ptr[t + 1];
  const long long n = e - b;
  const long long chunk = (n + 255) / 256;
  const long long lo = b + chunk * threadIdx.x;
  const long long hi = (lo + chunk < e) ? (lo + chunk) : e;
  double s = 0.0;
  for (long long k = lo; k < hi; k++) s += stage[g.src[k]];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) part[threadIdx.x] += part[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[g.target[t]] = part[0];
}

// ---------------------------------------------------------------------------------------------
// host launchers (called from plan.cu)
// ---------------------------------------------------------------------------------------------
static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template<int MODE>
static cudaError_t launch_dual_t(const FnDev& f, const EvalIO& io, cudaStream_t st) {
  const bool want_h = (MODE == FM_ALL) && !f.linear;
  const int nins = want_h ? f.n_total : f.n_primal;
  const int nslots = want_h ? f.slots_total : f.slots_primal;
  const size_t per_warp = (size_t) nslots * 32 * sizeof(Dual);
  const size_t tape_bytes = (size_t) nins * sizeof(Instr);
  int wpb = 4;
  while (wpb > 1 && tape_bytes + wpb * per_warp > 200 * 1024) wpb >>= 1;
  const size_t smem = tape_bytes + wpb * per_warp;
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(k_eval_dual<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
  if (e != cudaSuccess) return e;
  const int apw = (f.ir <= 32) ? (32 / f.ir) : 1;
  const long long napp = f.app_end - f.app_begin;
  if (napp <= 0) return cudaSuccess;
  const long long nitems = (napp + apw - 1) / apw;
  long long blocks = (nitems + wpb - 1) / wpb;
  const long long maxblocks = (long long) num_sms() * 16;
  if (blocks > maxblocks) blocks = maxblocks;
  k_eval_dual<MODE><<<(unsigned) blocks, wpb * 32, smem, st>>>(f, io);
  return cudaGetLastError();
}

cudaError_t launch_eval(const FnDev& f, const EvalIO& io, int mode, cudaStream_t st) {
  if (mode == FM_VALUES) {
    const size_t per_warp = (size_t) f.slots_primal * 32 * sizeof(Dual);
    const size_t tape_bytes = (size_t) f.n_primal * sizeof(Instr);
    int wpb = 4;
    while (wpb > 1 && tape_bytes + wpb * per_warp > 200 * 1024) wpb >>= 1;
    const size_t smem = tape_bytes + wpb * per_warp;
    if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(k_eval_values, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    const long long napp = f.app_end - f.app_begin;
    if (napp <= 0) return cudaSuccess;
    long long blocks = (napp + wpb * 32 - 1) / (wpb * 32);
    const long long maxblocks = (long long) num_sms() * 16;
    if (blocks > maxblocks) blocks = maxblocks;
    k_eval_values<<<(unsigned) blocks, wpb * 32, smem, st>>>(f, io);
    return cudaGetLastError();
  }
  switch (mode) {
    case FM_GRAD: return launch_dual_t<FM_GRAD>(f, io, st);
    case FM_JAC: return launch_dual_t<FM_JAC>(f, io, st);
    case FM_JAC_GRAD: return launch_dual_t<FM_JAC_GRAD>(f, io, st);
    case FM_ALL: return launch_dual_t<FM_ALL>(f, io, st);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_gather(const GatherDev& small, const GatherDev& large, const double* stage, double* out,
                          cudaStream_t st) {
  if (small.ntargets > 0) {
    k_gather_small<<<(small.ntargets + 255) / 256, 256, 0, st>>>(small, stage, out);
  }
  if (large.ntargets > 0) {
    k_gather_large<<<large.ntargets, 256, 0, st>>>(large, stage, out);
  }
  return cudaGetLastError();
}

}  // namespace ab
