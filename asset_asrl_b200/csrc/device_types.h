// Plain-old-data descriptors passed by value to the kernels (kernels.cu) by the plan (plan.cu).
#pragma once
#include <stdint.h>
#include <vector_types.h>

#include "tape.h"

namespace ab {

enum FnRole : int { ROLE_OBJ = 0, ROLE_EQ = 1, ROLE_IQ = 2 };

// What a launch computes, mirroring the per-function methods of the reference's solver interface
// (/root/reference/src/VectorFunctions/VectorFunctionTypeErasure/SolverInterfaceSpecs.h:41-92).
enum FnMode : int {
  FM_VALUES = 0,    // constraints() / objective()
  FM_GRAD = 1,      // constraints_adjointgradient() / objective_gradient(): values + J^T lambda, no KKT writes
  FM_JAC = 2,       // constraints_jacobian(): values + Jacobian into KKT
  FM_JAC_GRAD = 3,  // constraints_jacobian_adjointgradient()
  FM_ALL = 4        // constraints_jacobian_adjointgradient_adjointhessian() / objective_gradient_hessian()
};

// A linked device program (expr.hpp: Program) for one block size.
struct ProgDev {
  const uint4* code;     // 4 x 16-byte instructions per bundle
  const double* consts;  // shared constants, copied to shared memory by the block
  const double* tables;  // interpolation-table pool read by OP_TABIDX / OP_TABLOAD (null when the program has none)
  int nconst;
  int nslots;            // lane-private 8-byte slots per thread
  int nthreads;          // block size the operand offsets were linked for
  int nbundles;
  unsigned smem_bytes;
};

// One function (all its applications) as the device sees it.
struct FnDev {
  ProgDev prog;        // program of the evaluation mode being launched
  int ir, orows;
  int nappconst;
  int role;            // FnRole
  int linear;          // type-level linear: no Hessian triplets exist
  int aff_begin;       // applications [aff_begin, aff_end) use the affine index maps linked into the program
  int aff_end;
  int app_begin;       // first application evaluated by this device
  int app_end;         // one past the last
  int E;               // KKT triplets per application (depends on role / linear)
  const int* vindex;   // [nappl][ir]   global variable index of input i
  const int* cindex;   // [nappl][orows] constraint row (within its own eq / iq block)
  const double* appconst;  // [nappl][nappconst]
  const int* kkt_dest;     // [nappl][E]: >=0 CSR value slot written directly; <0: ~idx into stage
  const int* f_dest;       // [nappl][orows]: >=0 index into FX (eq / iq block); <0: ~idx into stage
  long long g_stage;       // gradient contributions go to stage[g_stage + app*ir + i]
  long long v_stage;       // objective value contributions go to stage[v_stage + app]
};

// cut set of a sharded plan, all output arrays (k_cut_all)
struct CutAll {
  double* arr[6];
  const int* target[6];
  const unsigned char* flag[6];
  long long off[7];  // entries of array a sit at buf[off[a] .. off[a + 1])
  int active[6];     // does the last evaluation's mode produce the array
  double* buf;
};

struct EvalIO {
  const double* X;
  const double* L;   // LE or LI for constraints (indexed by cindex); unused for objectives
  double obj_scale;
  double* kkt;       // CSR values
  double* fx;        // FXE or FXI block
  double* stage;     // staging buffer for every multi-source / reduced output
};

// Fixed-order segmented gather-sum: out[target[t]] = sum_{s in [ptr[t], ptr[t+1])} stage[src[s]]
struct GatherDev {
  int ntargets;
  int maxparts;  // high fan-in lists: ceil(largest fan-in / GATHER_PART) (two-stage fixed-order sum)
  const long long* ptr;
  const int* src;
  const int* target;
};

// sources summed by one block of the first stage of a high fan-in gather
constexpr int GATHER_PART = 2048;

}  // namespace ab
