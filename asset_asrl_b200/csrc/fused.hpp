// Owner-writes evaluation of the mesh interior: one thread per segment owns the CSR rows (and gradient entries) of its
// segment's first CS-1 nodes, evaluates every function application that touches them -- its own defect, integral,
// path constraints and splines, plus the neighbouring applications' contributions to those rows, recomputed instead
// of exchanged -- sums them in registers and writes every value once, in address order.
//
// Why: with one kernel per function a segment's 7 KB CSR region is written as ~50 short runs with
// holes and the node-shared entries detour through the staging buffer and a gather; partly written lines cost a
// read-modify-write at the memory side and the staging detour moves more bytes than the matrix itself.  Arithmetic is
// 7 % of the evaluation, so recomputing the neighbour's share is cheap.
//
// The reference's equivalent is the per-application loop with mutex-guarded += on shared entries
// (/root/reference/src/VectorFunctions/DenseFunctionBase.h:1276-1391, :1413-1466).
#pragma once
#include <string>
#include <vector>

#include "jit.hpp"
#include "nlp.hpp"

namespace ab {

// Array codes of the fused kernel's address tables.
enum FusedArr : int { FA_SCALE = -1, FA_X = 0, FA_LE = 1, FA_KKT = 2, FA_FXE = 3, FA_STAGE = 4, FA_APPC = 5, FA_LI = 6, FA_FXI = 7,
                      FA_PGX = 8, FA_AGX = 9 };

struct FusedPlan {
  bool ok = false;
  std::string why;        // why the problem has no fused interior (ok == false)
  int anchor = -1;        // function (index into Problem::funcs) whose applications are the threads
  int t_lo = 0, t_hi = 0; // anchor applications [t_lo, t_hi) are fused; rel = t - t_lo
  long long vb = 0, vs = 0;  // thread t owns variables [vb + (t - t_lo) * vs, + vs)
  long long kb = 0, ks = 0;  // ... and CSR slots [kb + (t - t_lo) * ks, + ks)
  struct Inst {
    int fi;        // function
    int m;         // applications of fi per thread step
    long long a0;  // application evaluated by thread t_lo; thread t evaluates a0 + m * (t - t_lo)
    bool home;     // this thread is the one that writes the application's constraint values / objective value
  };
  std::vector<Inst> insts;
  // per function: applications [ex_lo, ex_hi) are covered completely by fused threads and are skipped by the
  // per-function path (ex_lo == ex_hi: none)
  std::vector<int> ex_lo, ex_hi;
  // The programs: loads are entries of `fin` (element base for rel = 0, stride per rel), outputs entries of the
  // piece's `fout`.  The owned rows are cut into consecutive row blocks, one kernel each: a kernel then keeps fewer
  // values alive (the whole thread's work needs ~400, which leaves one block per SM) and stays under the code size at
  // which instruction fetch starts to cost; shared subexpressions are recomputed per piece.
  struct Piece {
    Program prog;
    std::vector<JitOut> fout;
  };
  std::vector<Piece> pieces;
  // Residual-only evaluations of the same threads (the modes PSIOPT's line search calls once to three times per iterate,
  // /root/reference/src/Solvers/PSIOPT.cpp:811-978): evalRHS (4: gradients, constraint values, objective value), evalOGC
  // (5), evalOCC (6), evalOBJ (7) -- the same DAG, fewer outputs, one piece each; empty when the mode has no fused program.
  std::vector<Piece> residual[8];
  std::vector<JitOut> fin;
  std::vector<long long> appc_offset;  // per function: offset of its per-application constants in the concatenated array
  int n_inst_home = 0;
};

// Finds the fused interior of an analysed problem and builds its program (evalKKT mode: values, Jacobian, gradients,
// Hessian).  Never throws for "not fusable": returns ok == false with the reason.
FusedPlan build_fused(const Problem& prob, int min_threads, int npieces = 0 /* 0: by program size */);

// PTX of one piece of the fused evaluation (entry "ab_fused_kkt<piece>"); fills the output table entries (emission order) and the statistics.
// reg_budget > 0 overrides the default number of values kept in registers (the plan lowers it when the assembler spills)
// mode: 0 (evalKKT, fp.pieces) or 4..7 (fp.residual[mode], entry "ab_fused_m<mode>_<piece>")
std::string fused_ptx(const FusedPlan& fp, int piece, std::vector<JitOut>* outs, JitStats* stats, int reg_budget = 0, int mode = 0);

}  // namespace ab
