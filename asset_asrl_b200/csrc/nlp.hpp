// Host-side NLP assembly: function list -> KKT triplets -> upper-triangular CSR pattern -> the
// segment-to-CSR destination maps and fixed-order gather lists the kernels consume.
//
// Restates, without Eigen, what the reference does once per transcription in
// NonLinearProgram::make_NLP / getMATSpace / analyzeSparsity
// (/root/reference/src/Solvers/NonLinearProgram.cpp:25-39, :111-175, :198-265, :267-344) and
// DenseFunctionBase::getKKTSpace (/root/reference/src/VectorFunctions/DenseFunctionBase.h:1097-1129).
// The canonical triplet order is the reference's Threads=1 order.
#pragma once
#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

#include "expr.hpp"

namespace ab {

struct FuncRec {
  int role = 1;  // FnRole: 0 objective, 1 equality, 2 inequality
  int tape = -1;
  int ir = 0, orows = 0;
  bool linear = false;
  bool unique_constraints = true;  // false for accumulations (PhaseIndexer.cpp:41-76)
  int nappl = 0;
  int nappconst = 0;
  std::vector<int> vindex;       // [nappl][ir]
  std::vector<int> cindex;       // [nappl][orows]  (row within the eq / iq block)
  std::vector<double> appconst;  // [nappl][nappconst]
  std::string name;

  int triplets_per_app() const {
    int h = linear ? 0 : ir * (ir + 1) / 2;
    int j = (role == 0) ? 0 : ir * orows;
    return h + j;
  }
};

struct GatherList {
  std::vector<long long> ptr;  // ntargets + 1
  std::vector<int> src;        // indices into the staging buffer, ascending per target
  std::vector<int> target;     // output index per target
  int ntargets() const { return (int) target.size(); }
};

struct Problem {
  std::vector<Tape> tapes;
  std::vector<FuncRec> funcs;  // in user order; evaluation order is objectives, equalities, inequalities
  int nvars = 0, neq = 0, niq = 0;

  // ---- filled by analyze() ----
  int kktdim = 0;
  long long n_user = 0;    // user triplets
  long long n_solver = 0;  // solver-owned entries (slack ones, primal diag, slack diag, eq pivots, iq pivots)
  std::vector<int> order;            // evaluation order of funcs
  std::vector<long long> trip_base;  // per func (indexed like funcs): first user triplet
  std::vector<long long> fdest_base; // per func: offset into f_dest
  std::vector<int> indptr, indices;  // CSR (upper triangle, row-major), bit-exact vs analyzeSparsity
  std::vector<int> kktloc;           // [n_user + n_solver] == reference KKTLocations
  std::vector<int> kkt_dest;         // [n_user] >=0 direct CSR slot, <0 ~stage index
  std::vector<int> f_dest;           // per constraint output
  // staging layout
  long long stage_size = 0;
  long long stage_solver = 0;              // solver coefficients live at stage[stage_solver + i]
  std::vector<long long> g_stage;          // per func
  std::vector<long long> v_stage;          // per objective func (else -1)
  // gather lists: [0] fan-in <= LARGE_FANIN, [1] the rest
  GatherList g_kkt[2], g_pgx[2], g_agx[2], g_fxe[2], g_fxi[2], g_val[2];
  bool analyzed = false;

  void analyze();
  long long nnz_kept = -1;  // set when the pattern arrays have been released (Plan::release_host_analysis)
  long long nnz() const { return nnz_kept >= 0 ? nnz_kept : (long long) indices.size(); }

  // ---- window of a single-phase problem (rank-local analysis, SURVEY.md section 8e / NonLinearProgram.cpp:267-344) ----
  // A rank that evaluates mesh segments [seg_lo, seg_hi) of one long phase needs the structure of those segments and of a
  // few neighbours only.  window() keeps, of every function, the applications whose *home segment* -- the segment of the
  // lowest mesh variable they read -- lies within `halo` segments of the range, with all indices still global (variables,
  // constraint rows, dimensions), so that analyze() yields, for the rows of the window's interior, exactly the rows of
  // the whole problem's KKT pattern at a cost proportional to the window.  Applications homed inside [seg_lo, seg_hi)
  // are *owned*: the rank evaluates those and nothing else.  For a function with several applications the home is taken
  // over the inputs that change from one application to the next (a mesh-spacing constraint reads the phase's first and
  // last time in every application; its home is the node in between).  Applications that read no such variable
  // (parameter-only functions) are kept everywhere and owned by the last rank.
  struct Window {
    bool on = false;
    long long mesh_vars = 0;  // variables [0, mesh_vars) belong to mesh nodes; the rest are parameters
    int seg_vars = 0;         // variables per segment
    long long nseg = 0, seg_lo = 0, seg_hi = 0;
    int halo = 0;
    bool first = false, last = false;     // the range starts / ends at a real mesh end
    std::vector<int> app_base;            // per function: application number (in the whole problem) of the first kept one
    std::vector<int> nappl_whole;         // per function: applications in the whole problem
    std::vector<int> own_lo, own_hi;      // per function: kept applications [own_lo, own_hi) are owned
    // Far variables: parameters and the mesh variables that every application of some function reads (the phase's first
    // and last time in the mesh-spacing constraints): their gradient entries and the KKT entries among them collect
    // contributions from every rank.
    std::vector<int> far_mesh_vars;  // sorted
    bool is_far(long long v) const {
      return v >= mesh_vars || std::binary_search(far_mesh_vars.begin(), far_mesh_vars.end(), (int) v);
    }
    // does this rank own the rows of (ordinary) mesh variable v?
    bool owns_var(long long v) const {
      if (v >= mesh_vars) return false;
      const long long sg = std::min(v / seg_vars, nseg - 1);
      return sg >= seg_lo && sg < seg_hi;
    }
  } win;
  // Restricts the function list to the window (before analyze()); throws if the kept or owned applications of a function
  // are not one contiguous range.
  void window(long long mesh_vars, int seg_vars, long long nseg, long long seg_lo, long long seg_hi, int halo);

  // Index ranges [lo, hi) of X (which = 0), LE (1) or LI (2) that the problem's functions read: what a rank has to receive
  // of the iterate each step.  The whole array for an ordinary problem; for a windowed problem the variables / multiplier
  // rows of the kept applications (own segments + halo + far variables), runs closer than `gap` elements joined into one
  // copy.  Ascending, disjoint.
  std::vector<std::pair<long long, long long>> input_ranges(int which, long long gap = 512) const;
};

constexpr int LARGE_FANIN = 96;

// ---------------------------------------------------------------------------------------------------------------
// One rank's share of the evaluation when a problem is split over several GPUs.  Every function's
// applications are cut into `world` contiguous ranges (the mirror of SolverIndexingData::thread_split,
// /root/reference/src/VectorFunctions/IndexingData.h:117-146); a rank evaluates its range, which fills the direct
// outputs of those applications and their staging slots.  An output with several sources is summed by the rank(s)
// holding its sources; outputs whose sources sit on more than one rank (the shared node at a cut, the control spline
// straddling it, parameter blocks, accumulated constraints, the objective value) form the *cut set*: each contributing
// rank sums its own sources, the partial sums are added across ranks (one all-reduce over the packed cut buffer) and the
// owner -- the lowest contributing rank -- keeps the result, so that every output ends up non-zero on exactly one rank.
// ---------------------------------------------------------------------------------------------------------------
enum OutArray : int { OA_KKT = 0, OA_PGX, OA_AGX, OA_FXE, OA_FXI, OA_VAL, OA_COUNT };
struct RankPartition {
  int rank = 0, world = 1;
  std::vector<int> app_begin, app_end;  // per function, indexed like Problem::funcs
  GatherList g[OA_COUNT][2];            // the problem's gather lists restricted to this rank's sources
  std::vector<int> cut_target[OA_COUNT];            // output index, ascending; identical on every rank
  std::vector<unsigned char> cut_owner[OA_COUNT];   // rank that keeps the summed value
  std::vector<unsigned char> cut_mine[OA_COUNT];    // 1 if this rank holds sources of the target
  long long cut_total() const {
    long long n = 0;
    for (int a = 0; a < OA_COUNT; a++) n += (long long) cut_target[a].size();
    return n;
  }
};
RankPartition make_partition(const Problem& prob, int rank, int world);

}  // namespace ab
