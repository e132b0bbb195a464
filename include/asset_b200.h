/*
 * asset_b200.h -- C ABI of the B200-native KKT evaluation path for ASSET-style collocation NLPs.
 *
 * Plain pointers and sizes only; every function returns 0 on success or a non-zero status
 * (CUDA error code for device failures, AB_ERR_* otherwise) and never throws across the boundary.
 * ab_last_error() returns the message of the last failure on the calling thread.
 *
 * Reference interfaces each group replaces (paths relative to the reference tree):
 *   ab_graph_*    the expression types behind vf.Arguments / VectorFunction
 *                 (src/VectorFunctions/VectorFunctionTypeErasure/GenericFunction.h:87-405,
 *                  src/VectorFunctions/CommonFunctions/ headers) -- recorded as a DAG, lowered to a tape.
 *   ab_phase_*    ODEPhase transcription: PhaseIndexer (src/OptimalControl/PhaseIndexer.h:63-99,
 *                 PhaseIndexer.cpp:132-410) and transcribe_phase (ODEPhaseBase.cpp:1364-1381).
 *   ab_problem_*  NonLinearProgram set-up: make_NLP / getMATSpace / analyzeSparsity
 *                 (src/Solvers/NonLinearProgram.cpp:25-39, :111-175, :267-344).
 *   ab_plan_*     the eight NonLinearProgram::eval* entry points PSIOPT calls
 *                 (src/Solvers/NonLinearProgram.h:411-478; PSIOPT.h:640-726), i.e. the per-function
 *                 SolverConstraintSpec / SolverObjectiveSpec methods
 *                 (src/VectorFunctions/VectorFunctionTypeErasure/SolverInterfaceSpecs.h:41-92, :254-277)
 *                 fused into device launches.
 */
#ifndef ASSET_B200_H
#define ASSET_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ab_graph ab_graph;
typedef struct ab_phase ab_phase;
typedef struct ab_problem ab_problem;
typedef struct ab_plan ab_plan;

enum { AB_OK = 0, AB_ERR_INVALID = -1, AB_ERR_RUNTIME = -2, AB_ERR_NO_DEVICE = -3 };

/* eval modes == NonLinearProgram::evalKKT / KKTNO / SOE / AUG / RHS / OGC / OCC / OBJ */
enum { AB_EVAL_KKT = 0, AB_EVAL_KKTNO = 1, AB_EVAL_SOE = 2, AB_EVAL_AUG = 3, AB_EVAL_RHS = 4, AB_EVAL_OGC = 5, AB_EVAL_OCC = 6, AB_EVAL_OBJ = 7 };
/* function roles */
enum { AB_ROLE_OBJECTIVE = 0, AB_ROLE_EQUALITY = 1, AB_ROLE_INEQUALITY = 2, AB_ROLE_INTEGRAL_OBJECTIVE = 3, AB_ROLE_INTEGRAL_PARAM = 4 };
/* device pointers handed to a device-resident consumer */
enum { AB_PTR_X = 0, AB_PTR_LE = 1, AB_PTR_LI = 2, AB_PTR_KKT = 3, AB_PTR_FXE = 4, AB_PTR_FXI = 5, AB_PTR_PGX = 6, AB_PTR_AGX = 7, AB_PTR_VAL = 8 };

/* expression node kinds (ab_graph_node); children / dparams / iparams per kind as listed */
enum {
  AB_VK_ARGS = 0,        /* Arguments(ir): identity */
  AB_VK_SEGMENT = 1,     /* child[start : start+size]        iparams = {start, size} */
  AB_VK_ELEMENTS = 2,    /* child[iparams...] */
  AB_VK_CONSTANT = 3,    /* constant vector dparams */
  AB_VK_STACK = 4,       /* concatenated outputs of the children */
  AB_VK_SUM = 5,         /* sum of the children */
  AB_VK_DIFF = 6,        /* child0 - child1 */
  AB_VK_SCALED = 7,      /* child0 * dparams[0] */
  AB_VK_ROWSCALED = 8,   /* child0 .* dparams */
  AB_VK_ADDVEC = 9,      /* child0 + dparams */
  AB_VK_VECMINUS = 10,   /* dparams - child0 */
  AB_VK_UNARY = 11,      /* cwise scalar op iparams[0] on child0 (dparams[0] = power for pow) */
  AB_VK_NORM = 12,       /* |child0|^p, iparams = {p} */
  AB_VK_NORMALIZED = 13, /* child0 / |child0|^p, iparams = {p} */
  AB_VK_DOT = 14, AB_VK_CROSS = 15, AB_VK_CWISEPROD = 16, AB_VK_CWISEQUOT = 17,
  AB_VK_VSPROD = 18,     /* vector child0 * scalar child1 */
  AB_VK_VSDIV = 19, AB_VK_CWISESUM = 20, AB_VK_ATAN2 = 21,
  AB_VK_NESTED = 22,     /* child0(child1(x)) */
  AB_VK_IFELSE = 23, AB_VK_COMPARE = 24, AB_VK_BOOL = 25, AB_VK_IOSCALED = 26, AB_VK_APPCONST = 27, AB_VK_MATVEC = 28,
  AB_VK_QUATPROD = 29,   /* Hamilton product child0 (x) child1, quaternions stored (x, y, z, w) (VectorProducts.h:28-30,229-236) */
  AB_VK_INTERP1D = 30,   /* InterpTable1D at child0 = t (InterpTable1D.h:9-483): iparams = {n, vlen, kind (0 cubic, 1 linear), even},
                            dparams = [ts (n), vs (vlen x n, row-major), dvs/dt (vlen x n; cubic only)] */
  AB_VK_INTERP2D = 31,   /* InterpTable2D at child0 = (x, y) (InterpTable2D.h): iparams = {nx, ny, kind, xeven, yeven},
                            dparams = [xs (nx), ys (ny), z, dz/dx, dz/dy, d2z/dxdy (ny x nx each, row-major; cubic: all four)] */
  AB_VK_MATPROD = 32,    /* matrix product child0 (r x k) * child1 (k x c), iparams = {r, k, c, child0 row-major, child1 row-major};
                            result column-major (MatrixFunctionProduct, CommonFunctions/MatrixProduct.h:19-120) */
  AB_VK_MATINV = 33,     /* inverse of the n x n matrix child0, iparams = {n, row-major} (CommonFunctions/MatrixInverse.h:6-58) */
  AB_VK_INTERPND = 34    /* InterpTable3D / InterpTable4D at child0 = the D coordinates (CommonFunctions/InterpTable3D.h:395-560,
                            InterpTable4D.h): iparams = {D, kind, n_0.., even_0..}, dparams = [axes, then 2^D arrays (cubic: array b
                            = derivative along the axes whose bit is set in b) or one (linear), shape (n_0..n_{D-1}), C order] */
};
/* PhaseRegionFlags and TranscriptionModes of the reference (src/OptimalControl/OptimalControlFlags.h) */
enum { AB_REGION_FRONT = 1, AB_REGION_BACK = 2, AB_REGION_FRONTANDBACK = 3, AB_REGION_BACKANDFRONT = 4, AB_REGION_PATH = 5,
       AB_REGION_INNERPATH = 6, AB_REGION_NODALPATH = 7, AB_REGION_PAIRWISEPATH = 9, AB_REGION_PARAMS = 12,
       AB_REGION_ODEPARAMS = 13, AB_REGION_STATICPARAMS = 14 };
enum { AB_TM_LGL3 = 0, AB_TM_LGL5 = 1, AB_TM_LGL7 = 2, AB_TM_TRAPEZOIDAL = 3 };

const char* ab_last_error(void);
int ab_version(void);
int ab_device_count(void);

/* ---------------- expression graph ---------------- */
ab_graph* ab_graph_create(void);
void ab_graph_destroy(ab_graph* g);
/* Adds a node and returns its id (>= 0) or a negative status.  `kind` is an ab::VKind; `tlin` = -1 lets the
 * library derive the type-level linearity from the reference's rule table, `raw` = 1 marks a value that is
 * not wrapped in a GenericFunction (Arguments / Segment / Element or a C++-built helper). */
int ab_graph_node(ab_graph* g, int kind, int ir, int orows, const int* children, int nchildren, const double* dparams,
                  int ndparams, const int* iparams, int niparams, int tlin, int raw);
int ab_graph_node_info(const ab_graph* g, int node, int* ir, int* orows, int* tlin, int* raw);
/* Writes an s-expression description of `node`; returns the full length (call again with a larger buffer). */
long long ab_graph_describe(const ab_graph* g, int node, char* buf, long long cap);

/* ---------------- problem: tapes + function list + sparsity ---------------- */
ab_problem* ab_problem_create(int nvars, int neq, int niq);
void ab_problem_destroy(ab_problem* p);
int ab_problem_set_dims(ab_problem* p, int nvars, int neq, int niq);
/* Lowers an expression to a device tape; returns the tape id. */
int ab_problem_add_tape(ab_problem* p, const ab_graph* g, int node, int with_adjoint);
/* Adds a function = (tape, Vindex, Cindex); vindex is [nappl][ir], cindex [nappl][orows] (ignored for objectives),
 * appconst [nappl][nappconst] or NULL.  Returns the function id. */
int ab_problem_add_function(ab_problem* p, int role, int tape, int linear, int unique_constraints, int nappl,
                            const int* vindex, const int* cindex, const double* appconst);
int ab_problem_analyze(ab_problem* p);
/* out[0..9] = nvars, neq, niq, kktdim, nnz, n_user_triplets, n_solver_entries, stage_size, nfunctions, ntapes */
int ab_problem_dims(const ab_problem* p, long long out[10]);
int ab_problem_pattern(const ab_problem* p, int* indptr, int* indices);
/* KKTLocations of every triplet in canonical (Threads = 1) order: user triplets then solver-owned entries. */
int ab_problem_kkt_locations(const ab_problem* p, int* out);
/* out[0..7] = role, tape, ir, orows, linear, unique_constraints, nappl, nappconst */
int ab_problem_function_info(const ab_problem* p, int f, int out[8]);
int ab_problem_function_indices(const ab_problem* p, int f, int* vindex, int* cindex);
int ab_problem_function_appconst(const ab_problem* p, int f, double* out /* [nappl][nappconst] */);
/* out[0..7] = n_instr, n_primal, slots_primal, slots_total, nin, nout, nappconst, has_adjoint */
int ab_problem_tape_info(const ab_problem* p, int t, int out[8]);
/* Text form of what tape `t` was lowered from ("expr <s-expression>" or a built-in transcription function with its
 * sizes); returns the full length.  Introspection for tests: the CPU oracle rebuilds the function from it. */
long long ab_problem_tape_describe(const ab_problem* p, int t, char* buf, long long cap);
/* Size of the device program of function `f` in an evaluation mode (0 values, 1 +gradient, 2 +Jacobian, 3 both, 4 all):
 * out[0..7] = operations, fused multiply-adds, special ops, loads, output stores, bundles, lane-private slots, constants */
int ab_problem_program_stats(const ab_problem* p, int f, int mode, long long out[8]);
/* Copies the (unlinked) program: 4 * bundles 16-byte instructions (asset_asrl_b200/csrc/expr.hpp: PInstr) and the constants. */
int ab_problem_program_code(const ab_problem* p, int f, int mode, void* code, double* consts);
/* The interpolation-table pool the program of function `f` reads (InterpTable1D / InterpTable2D data,
 * /root/reference/src/VectorFunctions/CommonFunctions/InterpTable1D.h:18-29, InterpTable2D.h:14-36): copies up to `cap`
 * doubles, returns the pool's length (0 for a function without tables) or a negative status. */
long long ab_problem_program_tables(const ab_problem* p, int f, int mode, double* out, long long cap);
/* PTX text of the specialised kernel the plan would assemble for function `f` in `mode` (asset_asrl_b200/csrc/jit.hpp):
 * `indexed` = 0 the affine variant run over the bulk of the mesh, 1 the indexed variant run over the mesh ends;
 * returns the full length, 0 if the function has no affine application range, or a negative status.
 * ab_problem_analyze() must have run.  ab_ptx_assemble assembles PTX for sm_100a in-process (no GPU needed):
 * out[0..2] = cubin bytes, registers per thread, spill bytes per thread. */
long long ab_problem_kernel_ptx(const ab_problem* p, int f, int mode, int indexed, char* buf, long long cap);
int ab_ptx_assemble(const char* ptx, long long out[3]);
/* copies n_instr 16-byte instructions (see asset_asrl_b200/csrc/tape.h) */
int ab_problem_tape_code(const ab_problem* p, int t, void* out);

/* ---------------- phase transcription ---------------- */
ab_phase* ab_phase_create(ab_graph* g, int ode_node, int xv, int uv, int pv, int transcription_mode, int nsegments);
void ab_phase_destroy(ab_phase* ph);
int ab_phase_set_control_mode(ab_phase* ph, int mode);
int ab_phase_set_integral_mode(ab_phase* ph, int mode);
int ab_phase_set_static_params(ab_phase* ph, int n);
int ab_phase_set_mesh(ab_phase* ph, int nbins, const double* bin_spacing /* nbins+1 */, const int* defs_per_bin);
int ab_phase_add_function(ab_phase* ph, int role, int node, int region, const int* xtuv, int nxtuv, const int* opv, int nopv,
                          const int* spv, int nspv, const int* ext, int next);
int ab_phase_clear_functions(ab_phase* ph);
/* out[0..7] = numPhaseVars, numStates, numDefects, cardinalStates, XtUVars, PVars, StaticPVars, blocked */
int ab_phase_dims(ab_phase* ph, int out[8]);
int ab_phase_state_spacing(ab_phase* ph, double* out /* numStates */);
/* Appends the phase's functions to `p` with variable / equality / inequality offsets; out = {numPhaseVars, numEq, numIq}. */
int ab_phase_transcribe(ab_phase* ph, ab_problem* p, int voff, int eoff, int ioff, int out[3]);
/* PhaseIndexer::makeSolverInput (PhaseIndexer.cpp:412-445): traj is [numStates][XtUVars + PVars]. */
/* Global variable indices of a phase region ([nappl][ir], out = {nappl, ir}); vindex == NULL only sizes.  Replaces
 * PhaseIndexer::make_Vindex_Cindex as called for link functions (src/OptimalControl/OptimalControlProblem.cpp:659-764).
 * Valid after ab_phase_transcribe (the phase's variable offset is fixed there). */
int ab_phase_region_vindex(ab_phase* ph, int region, const int* xtuv, int nxtuv, const int* opv, int nopv, const int* spv,
                           int nspv, int* vindex, int cap, int out[2]);
int ab_phase_make_input(ab_phase* ph, const double* traj, const double* static_params, double* X);

/* ---------------- device plan ---------------- */
/* Owner-writes fused interior of an analysed problem, host side only: out = {fusable, anchor
 * function, first fused segment, one past the last, variables per segment, CSR slots per segment, application families
 * per thread, of which home, distinct loads, outputs, FMAs over all pieces, number of pieces (kernels)}; `why` receives the
 * reason when the problem has no fused interior. */
long long ab_problem_fused_info(const ab_problem* p, int min_threads, long long out[12], char* why, long long cap);
/* Layout of the fused interior (introspection for tests/test_fused_flow_model.py): out = {first owned variable, variables
 * per thread, first owned CSR slot, slots per thread}; ex_lo / ex_hi (one entry per function, may be NULL): applications
 * [ex_lo, ex_hi) are evaluated by the fused threads and skipped by the per-function path. */
int ab_problem_fused_layout(const ab_problem* p, int min_threads, long long out[4], int* ex_lo, int* ex_hi);
/* Destination maps of function `f` after ab_problem_analyze (the device analogue of KKTLocations / the RHS slot starts,
 * src/Solvers/NonLinearProgram.cpp:318-335, src/VectorFunctions/IndexingData.h:96-115): per triplet and per constraint
 * output, >= 0 the slot of the output array written directly, < 0 the complement of a staging slot summed by the gather
 * lists; out = {first staging slot of the gradient contributions, of the objective-value contributions (-1: none)}. */
int ab_problem_dest_maps(const ab_problem* p, int f, int* kkt_dest, int* f_dest, long long out[2]);
/* PTX text of one piece of the fused evaluation (0 if there is none); buf == NULL only sizes. */
long long ab_problem_fused_ptx(const ab_problem* p, int min_threads, int piece, char* buf, long long cap);
/* ... of the residual-only fused kernel of an evaluation mode (AB_EVAL_RHS, _OGC, _OCC, _OBJ: the evaluations of PSIOPT's
 * line search, src/Solvers/PSIOPT.cpp:811-978); 0 if the mode has none. */
long long ab_problem_fused_residual_ptx(const ab_problem* p, int min_threads, int mode, char* buf, long long cap);
/* ... of piece `piece` of the fused kernels of any evaluation mode but evalKKT: the partial-matrix evaluations
 * AB_EVAL_KKTNO, _SOE, _AUG (src/Solvers/NonLinearProgram.cpp:540-666; same row blocks as the evalKKT pieces, each owned
 * matrix entry written with the contributions the mode computes) or a residual-only mode (piece 0); 0 if there is none. */
long long ab_problem_fused_mode_ptx(const ab_problem* p, int min_threads, int mode, int piece, char* buf, long long cap);
/* Rank-local analysis of one long phase (replaces the whole-problem pass of NonLinearProgram::analyzeSparsity,
 * /root/reference/src/Solvers/NonLinearProgram.cpp:267-344, by a pass over the rank's own segments).  Call between the
 * transcription and ab_problem_analyze: keeps, of every function, the applications whose lowest mesh variable lies in
 * segments [seg_lo - halo, seg_hi + halo) -- indices and dimensions stay those of the whole problem -- and marks those
 * homed in [seg_lo, seg_hi) as owned.  Variables [0, mesh_vars) belong to mesh nodes, seg_vars of them per segment.
 * A plan created from a windowed problem (ab_plan_create_sharded with the rank and world of the split) evaluates evalKKT
 * for the owned applications only: the CSR rows of the owned segments come out complete, in the windowed problem's own
 * pattern, whose rows inside the window are the whole problem's rows; entries that every rank contributes to (parameter
 * rows, accumulated constraints, the objective value) go through the cut buffer as for any sharded plan. */
int ab_problem_window(ab_problem* p, long long mesh_vars, int seg_vars, long long nseg, long long seg_lo, long long seg_hi, int halo);
/* out = {application number in the whole problem of the first kept one, applications in the whole problem, first owned
 * (kept numbering), one past the last owned} of function `fidx` */
int ab_problem_window_info(const ab_problem* p, int fidx, int out[4]);
/* mesh variables that every application of some function reads (entries among them are summed over the ranks); returns
 * their number and fills up to `cap` of them, ascending */
int ab_problem_window_far_vars(const ab_problem* p, int* out, int cap);
int ab_plan_create(const ab_problem* p, int device, ab_plan** out);
/* One rank's share of a problem split over `world` GPUs (the thread split of
 * SolverIndexingData::thread_split, src/VectorFunctions/IndexingData.h:117-146, across devices).  After ab_plan_eval:
 * ab_plan_cut_pack, then the caller sums ab_plan_cut_buffer() (ab_plan_cut_count() doubles, device memory) over all
 * ranks on the plan's stream (NCCL all-reduce), then ab_plan_cut_unpack.  Every output is then non-zero on exactly one
 * rank: the per-rank value arrays are disjoint slices whose sum is the reference's result. */
int ab_plan_create_sharded(const ab_problem* p, int device, int rank, int world, ab_plan** out);
/* Segments the fused owner-writes kernel evaluated in the last ab_plan_eval (0: the per-function path did everything). */
int ab_plan_fused_segments(const ab_plan* plan);
long long ab_plan_cut_count(const ab_plan* plan);
/* Index ranges [lo, hi) of output array `array` (0 KKT values, 1 PGX, 2 AGX, 3 FXE, 4 FXI, 5 val) that a sharded plan
 * can write; returns their number and fills up to `cap` pairs.  ab_plan_get_outputs copies only these ranges into the
 * caller's full-size arrays; ab_plan_get_outputs_packed copies them back to back into compact arrays. */
int ab_plan_output_ranges(const ab_plan* plan, int array, long long* out, int cap);
int ab_plan_get_outputs_packed(ab_plan* plan, double* kkt, double* pgx, double* agx, double* fxe, double* fxi, double* val);
void* ab_plan_cut_buffer(ab_plan* plan);
int ab_plan_cut_pack(ab_plan* plan);
int ab_plan_cut_unpack(ab_plan* plan);
void ab_plan_destroy(ab_plan* plan);

/* Host-side view of a rank's partition (application ranges, restricted gather lists, cut set); no GPU needed.
 * array: 0 KKT values, 1 PGX, 2 AGX, 3 FXE, 4 FXI, 5 val; level: 0 low fan-in list, 1 high fan-in list. */
typedef struct ab_partition ab_partition;
ab_partition* ab_partition_create(const ab_problem* p, int rank, int world);
void ab_partition_destroy(ab_partition* q);
int ab_partition_app_range(const ab_partition* q, int f, int out[2]);
int ab_partition_gather_sizes(const ab_partition* q, int array, int level, long long out[2] /* targets, sources */);
int ab_partition_gather(const ab_partition* q, int array, int level, long long* ptr, int* src, int* target);
long long ab_partition_cut_size(const ab_partition* q, int array);
int ab_partition_cut(const ab_partition* q, int array, int* target, int* owner, int* mine);
int ab_problem_gather_sizes(const ab_problem* p, int array, int level, long long out[3] /* targets, sources, stage size */);
int ab_problem_gather(const ab_problem* p, int array, int level, long long* ptr, int* src, int* target);
int ab_plan_set_inputs(ab_plan* plan, const double* X, const double* LE, const double* LI); /* host -> device */
/* The index ranges [lo, hi) of X (which = 0), LE (1), LI (2) the problem's functions read, as pairs in out[0 .. 2*cap);
 * returns the number of ranges.  The whole array for an ordinary problem; for a windowed problem (ab_problem_window) the
 * rank's own segments, its halo and the far variables.  The reference hands every thread the whole iterate
 * (/root/reference/src/Solvers/NonLinearProgram.cpp:473-537); a rank with its own memory needs only this slice of it.
 * ab_plan_input_ranges: the same for a plan (kept after a windowed plan has released its host analysis). */
long long ab_problem_input_ranges(const ab_problem* p, int which, long long* out, long long cap);
long long ab_plan_input_ranges(const ab_plan* plan, int which, long long* out, long long cap);
/* ab_plan_set_inputs that copies only those ranges of the (full-length) host arrays: the per-step host -> device traffic of
 * a rank follows its share of the phase, not the phase. */
int ab_plan_set_inputs_sliced(ab_plan* plan, const double* X, const double* LE, const double* LI);
int ab_plan_set_solver_coeffs(ab_plan* plan, const double* coeffs /* n_solver_entries */);
int ab_plan_eval(ab_plan* plan, int mode, double obj_scale); /* asynchronous on the plan's stream */
int ab_plan_sync(ab_plan* plan);
int ab_plan_get_outputs(ab_plan* plan, double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt_values);
/* host buffers in, host buffers out: H2D + eval + D2H, the call the reference's NonLinearProgram::eval* maps to */
int ab_plan_eval_host(ab_plan* plan, int mode, double obj_scale, const double* X, const double* LE, const double* LI,
                      double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt_values);
/* Per-launch device timing of the last eval (cudaEvent pairs on the plan's stream).  ab_plan_profile_read synchronises
 * the stream, writes up to `cap` durations in ms and returns the number of launches (or a negative status). */
int ab_plan_profile(ab_plan* plan, int enable); /* 0 off, 1 last eval only, 2 accumulate over evals */
int ab_plan_profile_read(ab_plan* plan, float* ms, int cap);
long long ab_plan_profile_label(ab_plan* plan, int k, char* buf, long long cap);
/* Checking aid: with ASSET_B200_GUARD=1 in the environment when the plan is created, every device allocation sits between
 * two guard bands; returns 0 if all are intact, k > 0 if the k-th allocation's band was written (ab_last_error says
 * where), a negative status if guard bands are off. */
int ab_plan_check_guards(ab_plan* plan);
void* ab_plan_device_ptr(ab_plan* plan, int which);
void* ab_plan_stream(ab_plan* plan);
long long ab_plan_launches(const ab_plan* plan);
/* how many of those were specialised (assembled at run time) kernels rather than interpreter launches */
long long ab_plan_specialised_launches(const ab_plan* plan);
/* per function k (evaluation order) and mode: out[0..3] = state (1 specialised, -1 interpreter, 0 not used yet),
 * registers per thread, spill bytes per thread, applications in the affine range */
int ab_plan_kernel_info(const ab_plan* plan, int k, int mode, int out[4]);

#ifdef __cplusplus
}
#endif
#endif
