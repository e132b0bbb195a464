// Device-resident evaluation plan: owns the index maps, tapes and value arrays on one GPU and runs the
// eight NonLinearProgram::eval* modes (/root/reference/src/Solvers/NonLinearProgram.h:411-478).
#pragma once
#include <cuda_runtime.h>

#include <array>
#include <string>
#include <vector>

#include "device_types.h"
#include "fused.hpp"
#include "jit.hpp"
#include "nlp.hpp"

namespace ab {

enum EvalMode : int {
  EM_KKT = 0,    // evalKKT   NonLinearProgram.cpp:473-537
  EM_KKTNO = 1,  // evalKKTNO :540-583  (no objective)
  EM_SOE = 2,    // evalSOE   :585-617  (constraints + Jacobian)
  EM_AUG = 3,    // evalAUG   :618-666
  EM_RHS = 4,    // evalRHS   :347-382
  EM_OGC = 5,    // evalOGC   :384-415
  EM_OCC = 6,    // evalOCC   :417-445
  EM_OBJ = 7     // evalOBJ   :447-466
};

struct DevGather {
  GatherDev dev[2] = {};
  void* bufs[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

struct Plan {
  Problem prob;
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string last_error;

  std::vector<FnDev> fns;  // in evaluation order
  std::vector<std::array<ProgDev, 5>> progs;  // per function, per FnMode (nthreads == 0: mode not generated)
  std::vector<void*> owned;  // every device allocation, freed in the destructor
  // Specialised kernels (jit.hpp): compiled on first use of a (function, mode) pair for the affine application range
  // of the function; the mesh ends and functions without such a range run on the interpreter.
  bool use_jit = true;    // ASSET_B200_JIT=0 switches the specialiser off (interpreter only)
  int jit_min_apps = 32;  // ASSET_B200_JIT_MIN: smallest affine range worth a specialised kernel
  std::vector<AffineMaps> affs;                        // per function
  std::vector<std::array<Program, 5>> host_progs;      // unlinked programs, kept for the specialiser
  // [0] affine variant (bulk of the mesh), [1] indexed variant (mesh ends, functions without an affine range)
  std::vector<std::array<JitKernel, 5>> jits[2];          // per function, per FnMode
  std::vector<std::array<signed char, 5>> jit_state[2];   // 0 not tried, 1 ready, -1 not specialisable
  std::string jit_log;                                    // why the last refused specialisation was refused
  int ensure_jit(size_t k, int fm, int variant);  // returns 1 ready, 0 use the interpreter, < 0 CUDA error (negated)
  cudaStream_t side = nullptr;  // the small launches (mesh ends, single applications) overlap the bulk kernels here
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // The bulk kernel of the heaviest function (the defect) is latency-bound at 4 warps per SM and leaves most of the
  // register file and the memory system idle; the bulk kernels of the other functions run beside it on this stream.
  cudaStream_t bulk2 = nullptr;
  cudaEvent_t ev_join2 = nullptr;
  int heavy_fn = -1;  // index into fns
  long long jit_launches = 0;                          // specialised-kernel launches of the last eval

  // Owner-writes fused interior (fused.hpp), used by evalKKT on a single device: the per-function path then covers only
  // the applications at the mesh ends and functions outside the periodic interior, and runs first; the fused kernel
  // runs last on the stream, so where both write a value (entries the end applications share with the first / last
  // fused segments) the fused result stands.
  bool use_fused = true;  // ASSET_B200_FUSED=0 switches it off
  FusedPlan fused;
  std::vector<JitKernel> fused_jits;  // one kernel per piece (row block)
  signed char fused_state = 0;  // 0 not tried, 1 ready, -1 unavailable
  DevGather gkF, gpF, gaF;      // gather lists without the targets the fused kernel writes
  // Fused evaluation: destination maps of the applications the per-function kernels still evaluate (the mesh ends),
  // with every CSR slot that a fused thread owns redirected to a trash slot of the staging buffer.  The fused kernels
  // and the mesh ends then never write the same entry, so nothing orders them: the ends run beside the bulk on a pool
  // of side streams and the fused kernels cover their whole segment range in one launch per piece.  Per function (in
  // evaluation order): device array of the rows of applications [0, lo) followed by [hi, nappl); kdF_hi_off = rows in
  // the first chunk.
  std::vector<int*> kdF;
  std::vector<int> kdF_lo, kdF_hi;
  static constexpr int SIDE_POOL = 4;
  cudaStream_t side_pool[SIDE_POOL] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_pool[SIDE_POOL] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_ends = nullptr;
  DevGather gkS;                // KKT entries that hold solver coefficients alone: written when those change, not per evaluation
  bool solver_entries_valid = false;
  DevGather gkE, gpE, gaE;      // ... of which the low fan-in targets no fused kernel feeds: summed on the side stream
  DevGather geF, giF, gvF;      // sharded plans: this rank's share of the constraint-row / objective-value lists
  // cut set of the fused evaluation on a sharded plan: every gather target fed by the fused kernels' staged outputs
  // (objective value, entries of rows no segment owns); rank 0 owns them all
  int* d_cutF_target[OA_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  unsigned char* d_cutF_flag[OA_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  long long cutF_n[OA_COUNT] = {0, 0, 0, 0, 0, 0};
  long long cutF_count = 0;
  // rank 0 of a sharded plan evaluates the applications at the far mesh end as well; what they write into entries that
  // the last rank's fused kernels own is cleared again on rank 0 (indices per output array)
  int* d_clear[OA_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  long long clear_n[OA_COUNT] = {0, 0, 0, 0, 0, 0};
  long long active_cut_count() const { return last_fused ? cutF_count : cut_count; }
  double* d_appc_all = nullptr; // per-application constants of all functions, concatenated (fused.appc_offset)
  int ensure_fused(int mode = 0);  // 1 ready, 0 unavailable, < 0 CUDA error (negated); mode 0 evalKKT, 4..7 residual-only modes
  std::vector<JitKernel> fused_res_jits[8];
  signed char fused_res_state[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  bool last_fused = false;      // the last eval ran the fused kernel

  double *d_X = nullptr, *d_LE = nullptr, *d_LI = nullptr;
  double *d_kkt = nullptr, *d_fxe = nullptr, *d_fxi = nullptr, *d_pgx = nullptr, *d_agx = nullptr, *d_val = nullptr;
  double* d_stage = nullptr;
  double* d_gscratch = nullptr;  // partial sums of the high fan-in gathers (one gather runs at a time)
  DevGather gk, gp, ga, ge, gi, gv;
  long long launches = 0;  // kernels launched by the last eval
  // optional per-launch device timing (cudaEvent pairs on `stream`), used by bench.py for the roofline of the
  // dominant kernel; labels are "<function name>" for tape launches and "gather:<output>" for the reductions
  int profiling = 0;  // 0 off, 1 keep the last eval's launches, 2 accumulate over evals until switched
  std::vector<cudaEvent_t> prof_events;
  std::vector<std::string> prof_labels;  // one per launch of the last eval
  int prof_begin(const std::string& label);
  int prof_end();
  int profile_read(float* ms, int cap);  // synchronises the stream; returns the number of launches

  ~Plan();
  // upload everything; returns 0 or a cudaError.  world > 1: this plan evaluates rank's share of every function
  // (RankPartition in nlp.hpp) and the cut set has to be reduced across ranks after each eval (cut_pack / all-reduce of
  // cut_buffer() by the caller / cut_unpack).
  int build(int device, int rank = 0, int world = 1);
  int rank = 0, world = 1;
  RankPartition part;
  int* d_cut_target[OA_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  unsigned char* d_cut_flag[OA_COUNT] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // bit0 mine, bit1 owner
  double* d_cutbuf = nullptr;
  long long cut_count = 0;
  // index ranges [lo, hi) of each output array this rank can write (direct outputs of its applications + its gather
  // targets, at a granularity of OUT_BLOCK elements); the read-back of a sharded plan copies only these ranges
  static constexpr long long OUT_BLOCK = 1 << 16;
  std::vector<std::array<long long, 2>> out_ranges[OA_COUNT];
  std::vector<char> out_blocks[OA_COUNT];  // touched blocks, collected while the plan is built
  void finalize_out_ranges();
  int last_mode = 0;
  int cut_pack();    // after eval(): partial sums of the cut set -> cut buffer (zeros where this rank has no sources)
  int cut_unpack();  // after the all-reduce: the owner keeps the sum, other contributors clear their partial
  // windowed plans (Problem::window): this rank's share of one long phase, analysed locally
  bool windowed = false;
  // what this plan's kernels read of X / LE / LI (Problem::input_ranges, taken before the analysis is released): a rank of
  // a windowed plan receives these ranges of the iterate each step, not the whole iterate
  std::vector<std::pair<long long, long long>> in_ranges[3];
  long long win_f_b = 0, win_f_e = 0;  // fused threads of the owned segments, relative to t_lo
  // Windowed plans: once everything is uploaded and the fused kernels are assembled, the plan's copy of the analysis
  // (pattern, triplet locations, destination maps, gather lists, index maps: ~20 KB per segment) is released -- eight
  // ranks of a 1.6 M-segment phase share one host.
  void release_host_analysis();
  bool release_windowed = false;
  void window_filter(GatherList (*fl)[2], GatherList (*flE)[2], std::vector<int>* cut_t, int true_world);
  int eval(int mode, double obj_scale);
  // eval() = enqueue() on the plan's streams, or -- from the second evaluation of a mode on, when every kernel it needs
  // has been built -- the replay of a CUDA graph captured from one enqueue(): the ~25 dependent launches and the
  // fork / join events of an evaluation then cost one graph launch on the host instead of one driver call each.
  int enqueue(int mode, double obj_scale);
  struct EvalGraph {
    cudaGraphExec_t exec = nullptr;
    double obj_scale = 0.0;
    long long launches = 0, jit_launches = 0;
    bool fused = false;
    int evals = 0;      // eager evaluations of this mode so far
    bool failed = false;
  };
  EvalGraph graphs[8];
  bool use_graphs = true;
  int set_solver_coeffs(const double* host_coeffs);

  template<class T>
  T* dalloc(size_t n);
  struct Guard {
    char* lo;
    char* hi;
    size_t hi_bytes;
  };
  std::vector<Guard> guards;  // ASSET_B200_GUARD=1: guard bands around every device allocation
  int check_guards();
  template<class T>
  T* dupload(const std::vector<T>& v);
};

// Largest application range of function `fi` over which every index map is affine in the application number.
AffineMaps function_affine(const Problem& prob, int fi);
cudaError_t launch_eval(const FnDev& f, cudaStream_t st, const EvalIO& io);
// pack: buf[i] = (active && mine) ? arr[target[i]] : 0;  unpack: owner: arr[target[i]] = buf[i], other contributors: 0
cudaError_t launch_cut_all(const CutAll& c, bool pack, cudaStream_t st);
cudaError_t launch_cut(bool pack, bool active, double* arr, const int* target, const unsigned char* flag, long long n,
                       double* buf, cudaStream_t st);
cudaError_t launch_clear(double* arr, const int* idx, long long n, cudaStream_t st);  // arr[idx[i]] = 0
cudaError_t launch_gather(const GatherDev& small, const GatherDev& large, const double* stage, double* out,
                          double* scratch /* large.ntargets * large.maxparts partial sums */, cudaStream_t st);

}  // namespace ab
