// Kernel specialiser: turns the scalar device program of one function (expr.hpp: Program) into a straight-line
// sm_100a kernel, written as PTX and assembled in-process with the static PTX compiler of the CUDA toolkit.
//
// Why: the program of a collocation defect is a few thousand fused multiply-adds on a few hundred live values
// (LGL5 two-body low thrust: 4302 DFMA, 867 outputs).  Interpreted, every DFMA costs three shared-memory loads, a
// store and ~25 integer instructions of decode; specialised, the values live in registers and the kernel is DFMA and
// output stores only, which is what lets the evaluation approach the HBM write roofline.  The C++ front end of nvcc
// / NVRTC needs minutes for one 5000-statement basic block, ptxas about a second, hence PTX and not CUDA C.
//
// A specialised kernel covers the *affine* application range of a function (all interior mesh segments, where every
// gather/scatter index is base + segment * stride, see AffineMaps); the few applications at the mesh ends and all
// single-application functions stay on the interpreter (kernels.cu), which reads explicit index maps.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "expr.hpp"

namespace ab {

struct JitOut;
struct JitSpec {
  std::string name;       // kernel entry name ([A-Za-z0-9_])
  int role = 1;           // FnRole
  int ir = 0, orows = 0;
  int nappconst = 0;
  int E = 0;              // KKT triplets per application
  // false: affine variant, addresses are base + rel*stride with the AffineMaps baked in as immediates (bulk of the
  // mesh); true: indexed variant, every gather/scatter index is read from the explicit maps (mesh ends, functions
  // without an affine range, single-application functions)
  bool indexed = false;
  long long g_stage = 0;  // gradient contributions go to stage[g_stage + app*ir + i]
  long long v_stage = 0;  // objective value contributions go to stage[v_stage + app]
  int nthreads = 128;     // block size (.maxntid)
  int reg_budget = 112;   // affine variant: values kept in registers; the rest are parked in shared memory
  // fused programs (fused.hpp): loads are entries of *fin, outputs (PO_TABLE) entries of *fout, both (array, element base
  // for rel = 0, stride per rel); array codes are FusedArr.  The multiplier array of inequalities travels in the
  // vindex parameter slot.
  const std::vector<JitOut>* fin = nullptr;
  const std::vector<JitOut>* fout = nullptr;
  bool lockstep = false;        // bar.sync at every output flush: the warps of a block share instruction-cache fills
  bool prefetch_table = true;   // output-table entries of the next flush are loaded one flush ahead
  bool flush_unroll4 = true;    // output walk in trips of four segments with rotating data / address registers
  bool skip_zero_outputs = true;  // structural zeros of the dense blocks are written once (plan build), not per eval
  int zero_run_min = 1;           // minimum length of a zero run enclosed by stored values to be skipped (sweep r1r: 1 is fastest)
  // Outputs leave straight from registers: runs of consecutive outputs are written with 32-byte vector stores on
  // 32-byte boundaries (a whole sector per lane; 16- and 8-byte stores only at the ends of a run), no shared-memory
  // transposition.  Measured (tools/direct_store_bench.cu, profiles/r2_sweeps/direct_store_r3a.txt): 6.2 TB/s at four
  // resident warps per SM against 1.2 TB/s for the tile walk at the same residency.  Which outputs share a sector depends
  // on where a thread's run starts; the start's phase is a function of (rel mod 4) known when the kernel is generated,
  // so every phase gets its own predicated store of the same registers.
  bool direct_stores = false;
  // Parked values are written and read back with volatile shared-memory accesses.  ptxas 12.9 was caught sinking a
  // plain st.shared of a recycled slot below the later ld.shared of the same address in a 6000-instruction block
  // (LGL7 fused piece; the reload then returns the slot's previous value): volatile accesses keep program order.
  bool volatile_parks = true;
  bool vector_loads = true;        // hoisted loads of consecutive inputs as ld.global.nc.v4 / .v2 where the alignment is thread-independent
  bool streaming_stores = false;   // direct stores carry the .cs (evict-first) cache operator
  // Direct stores only: finished outputs wait in a queue and leave spread evenly over the program's arithmetic instead
  // of in the bursts they are finished in (jit.cpp: drain_outputs); at most pace_burst outputs per arithmetic bundle.
  bool park_indexed = true;        // indexed variant: register budget with shared-memory parking like the affine variant (false: all values left to the assembler)
  bool phase_bases = true;         // one base register per (address group, alignment phase) for the 32-byte stores (jit.cpp: addr_sector)
  bool pace_outputs = false;
  int pace_burst = 2;
  int debug = 0;                // measurement only (wrong results): 1 no global stores, 2 no output walk at all
  bool hoist_loads = true;      // every input load is issued at the top of the kernel (one DRAM round trip, not 30)
};

struct JitStats {
  unsigned smem_bytes = 0;  // dynamic shared memory per block (transposition tile + parked values)
  int parked_slots = 0;     // lane-private shared-memory slots
  int parks = 0, reloads = 0;
  int reg_budget = 0;       // values kept in registers (as used)
};

// Kernel parameters, in order (see launch_jit).  begin/end: application range, relative to the first affine
// application am.a0 for the affine variant, absolute for the indexed variant.
struct JitArgs {
  const double* X;
  const double* L;
  double* kkt;
  double* fx;
  double* stage;
  const double* appconst;
  const int* vindex;
  const int* cindex;
  const int* kkt_dest;
  const int* f_dest;
  const long long* out_base;  // affine variant: absolute byte address of output slot k for rel = 0
  const int* out_stride;      // affine variant: byte stride of output slot k per application
  double obj_scale;
  int begin, end;
};

// Destination of one output of the affine variant, in emission order: element base + rel*stride of array `arr`
// (0 X, 1 L, 2 KKT values, 3 FX block, 4 staging buffer).  The plan turns these into the absolute address table.
struct JitOut {
  int arr;
  long long base, stride;
};

// Development knobs (ASSET_B200_JIT_REGS / _THREADS / _LOCKSTEP / _PREFETCH / _UNROLL / _HOIST) applied over the defaults.
void jit_env_knobs(JitSpec& sp);

// PTX text of the kernel; throws std::runtime_error when the function cannot be specialised (mixed direct/staged
// destinations inside the affine range, offsets out of range).
std::string emit_ptx(const Program& P, const JitSpec& spec, const AffineMaps& am /* unused when spec.indexed */,
                     std::vector<JitOut>* outs = nullptr, JitStats* stats = nullptr);

// Assembles PTX for sm_100a.  Works without a GPU (used by the CPU-side tests).  Returns false and fills `log` on error.
bool assemble_ptx(const std::string& ptx, std::vector<char>& cubin, std::string& log, int* nregs = nullptr,
                  int* spill_bytes = nullptr);

struct JitKernel {
  cudaLibrary_t lib = nullptr;
  cudaKernel_t kern = nullptr;
  int nthreads = 0;
  unsigned smem_bytes = 0;        // dynamic shared memory per block (set before load_jit)
  long long* out_base = nullptr;  // device tables of the affine variant (owned by the plan)
  int* out_stride = nullptr;
  int nregs = 0;        // from the assembler's report (0 if unknown)
  int spill_bytes = 0;  // spill stores + loads per thread
  bool ok() const { return kern != nullptr; }
};

cudaError_t load_jit(const std::vector<char>& cubin, const std::string& name, JitKernel& out);
void unload_jit(JitKernel& k);
cudaError_t launch_jit(const JitKernel& k, const JitArgs& a, cudaStream_t st);

}  // namespace ab
