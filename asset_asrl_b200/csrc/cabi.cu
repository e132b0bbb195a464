// extern "C" boundary (include/asset_b200.h).  No exceptions cross it.
#include <cuda_runtime.h>

#include <cstring>
#include <exception>
#include <stdexcept>
#include <algorithm>
#include <memory>
#include <string>

#include "../../include/asset_b200.h"
#include "phase.hpp"
#include "fused.hpp"
#include "plan.hpp"

using namespace ab;

struct ab_graph {
  Graph g;
};
struct ab_problem {
  Problem p;
  // the introspection entry points (ab_problem_fused_*) ask for the same fused plan again and again; keep the last one.
  // Keyed by everything build_fused reads besides the problem: min_threads, the function count / analysis state and the
  // ASSET_B200_FUSED_* switches.
  mutable std::shared_ptr<FusedPlan> fused_cache;
  mutable std::string fused_key;
};
static const FusedPlan& cached_fused(const ab_problem* p, int min_threads) {
  std::string key = std::to_string(min_threads) + "/" + std::to_string(p->p.funcs.size()) + "/" + std::to_string(p->p.nnz()) + "/"
                    + std::to_string((int) p->p.analyzed) + "/" + std::to_string((int) p->p.win.on);
  for (const char* v: {"ASSET_B200_FUSED_BLOCKED", "ASSET_B200_FUSED_PIECES", "ASSET_B200_FUSED_SLOTS", "ASSET_B200_FUSED_EVEN_SPLIT",
                       "ASSET_B200_FUSED_BALANCE", "ASSET_B200_FUSED_DIRECT", "ASSET_B200_FUSED_REGS", "ASSET_B200_FUSED_REGS_BY_PIECE",
                       "ASSET_B200_FUSED_ZRUN", "ASSET_B200_FUSED_THREADS", "ASSET_B200_JIT_REGS", "ASSET_B200_JIT_DEBUG"}) {
    const char* e = getenv(v);
    key += std::string("/") + (e ? e : "-");
  }
  if (!p->fused_cache || p->fused_key != key) {
    p->fused_cache = std::make_shared<FusedPlan>(build_fused(p->p, min_threads));
    p->fused_key = key;
  }
  return *p->fused_cache;
}
struct ab_phase {
  Phase ph;
};
struct ab_plan {
  Plan plan;
};

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define AB_TRY try {
#define AB_CATCH                                                     \
  }                                                                  \
  catch (const std::exception& e) { return fail(AB_ERR_INVALID, e.what()); } \
  catch (...) { return fail(AB_ERR_RUNTIME, "unknown exception"); }

extern "C" {

const char* ab_last_error(void) { return g_err.c_str(); }
int ab_version(void) { return 100; }
int ab_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// ------------------------------------------------------------------ graph
ab_graph* ab_graph_create(void) { return new (std::nothrow) ab_graph(); }
void ab_graph_destroy(ab_graph* g) { delete g; }

static bool auto_tlin(const Graph& g, const VNode& n) {
  auto all_seen = [&]() {
    for (int c: n.ch)
      if (!g.seen_lin(c)) return false;
    return true;
  };
  switch (n.kind) {
    case VK_ARGS:
    case VK_CONSTANT:
    case VK_APPCONST: return true;
    case VK_SEGMENT:
    case VK_ELEMENTS:
    case VK_STACK:
    case VK_SUM:
    case VK_DIFF:
    case VK_SCALED:
    case VK_ROWSCALED:
    case VK_ADDVEC:
    case VK_VECMINUS:
    case VK_CWISESUM:
    case VK_IOSCALED:
    case VK_MATVEC:
    case VK_NESTED: return all_seen();
    default: return false;
  }
}

int ab_graph_node(ab_graph* g, int kind, int ir, int orows, const int* ch, int nch, const double* d, int nd, const int* ip,
                  int nip, int tlin, int raw) {
  AB_TRY
  if (!g) return fail(AB_ERR_INVALID, "null graph");
  VNode n;
  n.kind = kind;
  n.ir = ir;
  n.orows = orows;
  n.ch.assign(ch, ch + nch);
  n.d.assign(d, d + nd);
  n.ip.assign(ip, ip + nip);
  for (int c: n.ch)
    if (c < 0 || c >= (int) g->g.nodes.size()) return fail(AB_ERR_INVALID, "child id out of range");
  n.raw = raw != 0;
  n.tlin = (tlin < 0) ? auto_tlin(g->g, n) : (tlin != 0);
  // structural validation by a trial lowering on a scratch DAG
  int id = g->g.add(n);
  try {
    SDag dag;
    std::vector<int> in(ir);
    for (int i = 0; i < ir; i++) in[i] = dag.input(i);
    lower(g->g, id, in, dag);
  } catch (...) {
    g->g.nodes.pop_back();
    throw;
  }
  return id;
  AB_CATCH
}

int ab_graph_node_info(const ab_graph* g, int node, int* ir, int* orows, int* tlin, int* raw) {
  AB_TRY
  const VNode& n = g->g.at(node);
  if (ir) *ir = n.ir;
  if (orows) *orows = n.orows;
  if (tlin) *tlin = n.tlin;
  if (raw) *raw = n.raw;
  return AB_OK;
  AB_CATCH
}

long long ab_graph_describe(const ab_graph* g, int node, char* buf, long long cap) {
  try {
    std::string s = g->g.describe(node);
    if (buf && cap > 0) {
      long long n = std::min<long long>(cap - 1, (long long) s.size());
      std::memcpy(buf, s.data(), (size_t) n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}

// ------------------------------------------------------------------ problem
ab_problem* ab_problem_create(int nvars, int neq, int niq) {
  ab_problem* p = new (std::nothrow) ab_problem();
  if (p) {
    p->p.nvars = nvars;
    p->p.neq = neq;
    p->p.niq = niq;
  }
  return p;
}
void ab_problem_destroy(ab_problem* p) { delete p; }
int ab_problem_set_dims(ab_problem* p, int nvars, int neq, int niq) {
  if (!p) return fail(AB_ERR_INVALID, "null problem");
  p->p.nvars = nvars;
  p->p.neq = neq;
  p->p.niq = niq;
  p->p.analyzed = false;
  return AB_OK;
}

int ab_problem_add_tape(ab_problem* p, const ab_graph* g, int node, int with_adjoint) {
  AB_TRY
  ScalarFn f = lower_function(g->g, node);
  if (with_adjoint) f.make_gradient();
  p->p.tapes.push_back(emit_tape(f, with_adjoint != 0));
  p->p.tapes.back().desc = "expr " + g->g.describe(node);
  p->p.tapes.back().fn = std::make_shared<ScalarFn>(f);
  return (int) p->p.tapes.size() - 1;
  AB_CATCH
}

int ab_problem_add_function(ab_problem* p, int role, int tape, int linear, int unique, int nappl, const int* vindex,
                            const int* cindex, const double* appconst) {
  AB_TRY
  if (tape < 0 || tape >= (int) p->p.tapes.size()) return fail(AB_ERR_INVALID, "tape id out of range");
  if (role < 0 || role > 2) return fail(AB_ERR_INVALID, "role must be 0 (objective), 1 (equality) or 2 (inequality)");
  const Tape& t = p->p.tapes[tape];
  FuncRec fr;
  fr.role = role;
  fr.tape = tape;
  fr.ir = t.nin;
  fr.orows = t.nout;
  fr.linear = linear != 0;
  fr.unique_constraints = unique != 0;
  fr.nappl = nappl;
  fr.nappconst = t.nappconst;
  fr.vindex.assign(vindex, vindex + (size_t) nappl * fr.ir);
  if (role != 0) fr.cindex.assign(cindex, cindex + (size_t) nappl * fr.orows);
  if (fr.nappconst > 0) {
    if (!appconst) return fail(AB_ERR_INVALID, "tape needs per-application constants");
    fr.appconst.assign(appconst, appconst + (size_t) nappl * fr.nappconst);
  }
  fr.name = "function" + std::to_string(p->p.funcs.size());
  p->p.funcs.push_back(std::move(fr));
  p->p.analyzed = false;
  return (int) p->p.funcs.size() - 1;
  AB_CATCH
}

int ab_problem_window(ab_problem* p, long long mesh_vars, int seg_vars, long long nseg, long long seg_lo, long long seg_hi, int halo) {
  AB_TRY
  p->p.window(mesh_vars, seg_vars, nseg, seg_lo, seg_hi, halo);
  return AB_OK;
  AB_CATCH
}
int ab_problem_window_info(const ab_problem* p, int fidx, int out[4]) {
  AB_TRY
  const Problem::Window& W = p->p.win;
  if (!W.on) return fail(AB_ERR_INVALID, "not a windowed problem");
  if (fidx < 0 || fidx >= (int) W.app_base.size()) return fail(AB_ERR_INVALID, "bad function index");
  out[0] = W.app_base[(size_t) fidx];
  out[1] = W.nappl_whole[(size_t) fidx];
  out[2] = W.own_lo[(size_t) fidx];
  out[3] = W.own_hi[(size_t) fidx];
  return AB_OK;
  AB_CATCH
}

int ab_problem_window_far_vars(const ab_problem* p, int* out, int cap) {
  const auto& v = p->p.win.far_mesh_vars;
  for (size_t k = 0; k < v.size() && (int) k < cap; k++) out[k] = v[k];
  return (int) v.size();
}

long long ab_problem_input_ranges(const ab_problem* p, int which, long long* out, long long cap) {
  try {
    const auto r = p->p.input_ranges(which);
    for (size_t k = 0; k < r.size() && (long long) k < cap; k++) {
      out[2 * k] = r[k].first;
      out[2 * k + 1] = r[k].second;
    }
    return (long long) r.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}

int ab_problem_analyze(ab_problem* p) {
  AB_TRY
  p->p.analyze();
  return AB_OK;
  AB_CATCH
}

int ab_problem_dims(const ab_problem* p, long long out[10]) {
  const Problem& q = p->p;
  out[0] = q.nvars;
  out[1] = q.neq;
  out[2] = q.niq;
  out[3] = q.kktdim;
  out[4] = q.nnz();
  out[5] = q.n_user;
  out[6] = q.n_solver;
  out[7] = q.stage_size;
  out[8] = (long long) q.funcs.size();
  out[9] = (long long) q.tapes.size();
  return AB_OK;
}

int ab_problem_pattern(const ab_problem* p, int* indptr, int* indices) {
  if (!p->p.analyzed) return fail(AB_ERR_INVALID, "problem not analyzed");
  std::memcpy(indptr, p->p.indptr.data(), p->p.indptr.size() * sizeof(int));
  std::memcpy(indices, p->p.indices.data(), p->p.indices.size() * sizeof(int));
  return AB_OK;
}
int ab_problem_kkt_locations(const ab_problem* p, int* out) {
  if (!p->p.analyzed) return fail(AB_ERR_INVALID, "problem not analyzed");
  std::memcpy(out, p->p.kktloc.data(), p->p.kktloc.size() * sizeof(int));
  return AB_OK;
}
int ab_problem_function_info(const ab_problem* p, int f, int out[8]) {
  AB_TRY
  const FuncRec& fr = p->p.funcs.at(f);
  out[0] = fr.role;
  out[1] = fr.tape;
  out[2] = fr.ir;
  out[3] = fr.orows;
  out[4] = fr.linear;
  out[5] = fr.unique_constraints;
  out[6] = fr.nappl;
  out[7] = fr.nappconst;
  return AB_OK;
  AB_CATCH
}
int ab_problem_function_indices(const ab_problem* p, int f, int* vindex, int* cindex) {
  AB_TRY
  const FuncRec& fr = p->p.funcs.at(f);
  if (vindex) std::memcpy(vindex, fr.vindex.data(), fr.vindex.size() * sizeof(int));
  if (cindex) std::memcpy(cindex, fr.cindex.data(), fr.cindex.size() * sizeof(int));
  return AB_OK;
  AB_CATCH
}
int ab_problem_function_appconst(const ab_problem* p, int f, double* out) {
  AB_TRY
  const FuncRec& fr = p->p.funcs.at(f);
  if (!fr.appconst.empty()) std::memcpy(out, fr.appconst.data(), fr.appconst.size() * sizeof(double));
  return AB_OK;
  AB_CATCH
}
int ab_problem_tape_info(const ab_problem* p, int t, int out[8]) {
  AB_TRY
  const Tape& tp = p->p.tapes.at(t);
  out[0] = (int) tp.code.size();
  out[1] = tp.n_primal;
  out[2] = tp.slots_primal;
  out[3] = tp.slots_total;
  out[4] = tp.nin;
  out[5] = tp.nout;
  out[6] = tp.nappconst;
  out[7] = tp.has_adjoint;
  return AB_OK;
  AB_CATCH
}
long long ab_problem_tape_describe(const ab_problem* p, int t, char* buf, long long cap) {
  try {
    const std::string& s = p->p.tapes.at(t).desc;
    if (buf && cap > 0) {
      long long n = std::min<long long>(cap - 1, (long long) s.size());
      std::memcpy(buf, s.data(), (size_t) n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}
static Program program_of(const ab_problem* p, int fidx, int mode) {
  const FuncRec& fr = p->p.funcs.at(fidx);
  const Tape& tp = p->p.tapes.at(fr.tape);
  if (!tp.fn) throw std::invalid_argument("tape has no lowered function");
  if (mode < 0 || mode > 4) throw std::invalid_argument("mode must be 0..4");
  ProgramSpec sp;
  sp.is_objective = fr.role == 0;
  sp.linear = fr.linear;
  sp.want_g = (mode == FM_GRAD || mode == FM_JAC_GRAD || mode == FM_ALL);
  sp.want_j = (mode >= FM_JAC);
  sp.want_h = (mode == FM_ALL);
  ScalarFn copy = *tp.fn;
  return emit_program(copy, sp);
}
int ab_problem_program_code(const ab_problem* p, int fidx, int mode, void* code, double* consts) {
  AB_TRY
  Program P = program_of(p, fidx, mode);
  if (code) std::memcpy(code, P.code.data(), P.code.size() * sizeof(PInstr));
  if (consts) std::memcpy(consts, P.consts.data(), P.consts.size() * sizeof(double));
  return AB_OK;
  AB_CATCH
}
long long ab_problem_program_tables(const ab_problem* p, int fidx, int mode, double* out, long long cap) {
  try {
    Program P = program_of(p, fidx, mode);
    const long long n = (long long) P.tables.size();
    if (out && cap > 0) std::memcpy(out, P.tables.data(), (size_t) std::min(n, cap) * sizeof(double));
    return n;
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}
int ab_problem_program_stats(const ab_problem* p, int fidx, int mode, long long out[8]) {
  AB_TRY
  Program P = program_of(p, fidx, mode);
  out[0] = P.n_ops;
  out[1] = P.n_fma;
  out[2] = P.n_special;
  out[3] = P.n_load;
  out[4] = P.n_out;
  out[5] = P.bundles();
  out[6] = P.nslots;
  out[7] = (long long) P.consts.size();
  return AB_OK;
  AB_CATCH
}
long long ab_problem_kernel_ptx(const ab_problem* p, int fidx, int mode, int indexed, char* buf, long long cap) {
  try {
    if (!p->p.analyzed) throw std::invalid_argument("ab_problem_analyze() first");
    const FuncRec& fr = p->p.funcs.at(fidx);
    Program P = program_of(p, fidx, mode);
    const AffineMaps am = function_affine(p->p, fidx);
    if (!indexed && am.a1 <= am.a0) return 0;
    JitSpec sp;
    sp.name = std::string(indexed ? "ab_idx_f" : "ab_seg_f") + std::to_string(fidx) + "_m" + std::to_string(mode);
    sp.indexed = indexed != 0;
    sp.E = fr.triplets_per_app();
    sp.role = fr.role;
    sp.ir = fr.ir;
    sp.orows = fr.orows;
    sp.nappconst = fr.nappconst;
    sp.g_stage = p->p.g_stage.at(fidx);
    sp.v_stage = p->p.v_stage.at(fidx);
    sp.nthreads = 64;
    jit_env_knobs(sp);
    const std::string s = emit_ptx(P, sp, am);
    if (buf && cap > 0) {
      const size_t n = std::min<size_t>(s.size(), (size_t) cap - 1);
      std::memcpy(buf, s.data(), n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}
int ab_ptx_assemble(const char* ptx, long long out[3]) {
  AB_TRY
  std::vector<char> cubin;
  std::string log;
  int nregs = 0, spill = 0;
  if (!assemble_ptx(ptx, cubin, log, &nregs, &spill)) throw std::runtime_error(log);
  out[0] = (long long) cubin.size();
  out[1] = nregs;
  out[2] = spill;
  return AB_OK;
  AB_CATCH
}
int ab_problem_tape_code(const ab_problem* p, int t, void* out) {
  AB_TRY
  const Tape& tp = p->p.tapes.at(t);
  std::memcpy(out, tp.code.data(), tp.code.size() * sizeof(Instr));
  return AB_OK;
  AB_CATCH
}

// ------------------------------------------------------------------ phase
ab_phase* ab_phase_create(ab_graph* g, int ode_node, int xv, int uv, int pv, int tmode, int nseg) {
  try {
    if (!g) throw std::invalid_argument("null graph");
    const VNode& n = g->g.at(ode_node);
    if (n.ir != xv + 1 + uv + pv || n.orows != xv) throw std::invalid_argument("ODE expression must map (XV+1+UV+PV) -> XV");
    if (tmode < TM_LGL3 || tmode > TM_TRAPEZOIDAL) throw std::invalid_argument("Invalid Transcription Method");
    if (nseg < 1) throw std::invalid_argument("number of segments must be positive");
    ab_phase* ph = new ab_phase();
    ph->ph.g = &g->g;
    ph->ph.ode_root = ode_node;
    ph->ph.XV = xv;
    ph->ph.UV = uv;
    ph->ph.PV = pv;
    ph->ph.tmode = tmode;
    ph->ph.defs_per_bin = {nseg};
    ph->ph.init_indexing();
    return ph;
  } catch (const std::exception& e) {
    fail(AB_ERR_INVALID, e.what());
    return nullptr;
  }
}
void ab_phase_destroy(ab_phase* ph) { delete ph; }
int ab_phase_set_control_mode(ab_phase* ph, int m) {
  if (m < 0 || m > 3) return fail(AB_ERR_INVALID, "bad control mode");
  ph->ph.cmode = m;
  ph->ph.init_indexing();
  return AB_OK;
}
int ab_phase_set_integral_mode(ab_phase* ph, int m) {
  if (m < 0 || m > 2) return fail(AB_ERR_INVALID, "bad integral mode");
  ph->ph.imode = m;
  return AB_OK;
}
int ab_phase_set_static_params(ab_phase* ph, int n) {
  if (n < 0) return fail(AB_ERR_INVALID, "negative static parameter count");
  ph->ph.SPV = n;
  ph->ph.init_indexing();
  return AB_OK;
}
int ab_phase_set_mesh(ab_phase* ph, int nbins, const double* sp, const int* defs) {
  AB_TRY
  if (nbins < 1) return fail(AB_ERR_INVALID, "need at least one bin");
  ph->ph.bin_spacing.assign(sp, sp + nbins + 1);
  ph->ph.defs_per_bin.assign(defs, defs + nbins);
  ph->ph.init_indexing();
  return AB_OK;
  AB_CATCH
}
int ab_phase_add_function(ab_phase* ph, int role, int node, int region, const int* x, int nx, const int* o, int no, const int* s,
                          int ns, const int* e, int ne) {
  AB_TRY
  StateFn sf;
  sf.root = node;
  sf.region = region;
  sf.xtuv.assign(x, x + nx);
  sf.opv.assign(o, o + no);
  sf.spv.assign(s, s + ns);
  sf.ext.assign(e, e + ne);
  ph->ph.g->at(node);
  if (region == R_ODEPARAMS || region == R_STATICPARAMS) sf.region = R_PARAMS;
  switch (role) {
    case AB_ROLE_OBJECTIVE: ph->ph.objs.push_back(sf); return (int) ph->ph.objs.size() - 1;
    case AB_ROLE_EQUALITY: ph->ph.eqs.push_back(sf); return (int) ph->ph.eqs.size() - 1;
    case AB_ROLE_INEQUALITY: ph->ph.iqs.push_back(sf); return (int) ph->ph.iqs.size() - 1;
    case AB_ROLE_INTEGRAL_OBJECTIVE: ph->ph.integrals.push_back(sf); return (int) ph->ph.integrals.size() - 1;
    case AB_ROLE_INTEGRAL_PARAM: ph->ph.param_integrals.push_back(sf); return (int) ph->ph.param_integrals.size() - 1;
    default: return fail(AB_ERR_INVALID, "bad role");
  }
  AB_CATCH
}
int ab_phase_clear_functions(ab_phase* ph) {
  ph->ph.eqs.clear();
  ph->ph.iqs.clear();
  ph->ph.objs.clear();
  ph->ph.integrals.clear();
  ph->ph.param_integrals.clear();
  return AB_OK;
}
int ab_phase_dims(ab_phase* ph, int out[8]) {
  AB_TRY
  ph->ph.init_indexing();
  const PhaseIndexer& ix = ph->ph.idx;
  out[0] = ix.numPhaseVars;
  out[1] = ix.numStates;
  out[2] = ix.numDefects;
  out[3] = ix.DCS;
  out[4] = ix.XtUV();
  out[5] = ix.PV;
  out[6] = ix.SPV;
  out[7] = ix.blocked;
  return AB_OK;
  AB_CATCH
}
int ab_phase_state_spacing(ab_phase* ph, double* out) {
  AB_TRY
  auto s = ph->ph.state_spacing();
  std::memcpy(out, s.data(), s.size() * sizeof(double));
  return AB_OK;
  AB_CATCH
}
int ab_phase_transcribe(ab_phase* ph, ab_problem* p, int voff, int eoff, int ioff, int out[3]) {
  AB_TRY
  ph->ph.transcribe(p->p, voff, eoff, ioff);
  out[0] = ph->ph.idx.numPhaseVars;
  out[1] = ph->ph.idx.numEq;
  out[2] = ph->ph.idx.numIq;
  p->p.analyzed = false;
  return AB_OK;
  AB_CATCH
}
int ab_phase_region_vindex(ab_phase* ph, int region, const int* xtuv, int nxtuv, const int* opv, int nopv, const int* spv,
                           int nspv, int* vindex, int cap, int out[2]) {
  AB_TRY
  // PhaseIndexer::make_Vindex_Cindex(reg, xtv, opv, spv, 1)[0] as OptimalControlProblem::make_link_Vindex_Cindex calls it
  // (OptimalControlProblem.cpp:688-689, :745-746); valid after ab_phase_transcribe, which fixed the phase's offsets.
  std::vector<int> x(xtuv, xtuv + nxtuv), o(opv, opv + nopv), s(spv, spv + nspv), V, Cx;
  int next = 0, nappl = 0;
  ph->ph.idx.make_vindex_cindex(region, x, o, s, 0, next, V, Cx, nappl);
  out[0] = nappl;
  out[1] = nappl > 0 ? (int) (V.size() / (size_t) nappl) : 0;
  if (vindex) {
    if ((size_t) cap < V.size()) return fail(AB_ERR_INVALID, "ab_phase_region_vindex: buffer too small");
    std::memcpy(vindex, V.data(), V.size() * sizeof(int));
  }
  return AB_OK;
  AB_CATCH
}
int ab_phase_make_input(ab_phase* ph, const double* traj, const double* sp, double* X) {
  AB_TRY
  ph->ph.init_indexing();
  const PhaseIndexer& ix = ph->ph.idx;
  const int n = ix.XtUV(), w = n + ix.PV, xt = ix.XtV();
  if (ix.blocked) {
    for (int i = 0; i < ix.numStates; i++)
      for (int k = 0; k < xt; k++) X[(size_t) i * xt + k] = traj[(size_t) i * w + k];
    for (int i = 0; i < ix.numDefects; i++)
      for (int k = 0; k < ix.UV; k++)
        X[(size_t) ix.numStates * xt + (size_t) i * ix.UV + k] = traj[(size_t) i * (ix.DCS - 1) * w + xt + k];
    for (int k = 0; k < ix.PV; k++) X[(size_t) ix.numStates * xt + (size_t) ix.numDefects * ix.UV + k] = traj[n + k];
  } else {
    for (int i = 0; i < ix.numStates; i++)
      for (int k = 0; k < n; k++) X[(size_t) i * n + k] = traj[(size_t) i * w + k];
    for (int k = 0; k < ix.PV; k++) X[(size_t) ix.numStates * n + k] = traj[n + k];
  }
  for (int k = 0; k < ix.SPV; k++) X[(size_t) ix.numPhaseVars - ix.SPV + k] = sp[k];
  return AB_OK;
  AB_CATCH
}

// ------------------------------------------------------------------ fused interior (host side; no GPU needed)
long long ab_problem_fused_info(const ab_problem* p, int min_threads, long long out[12], char* why, long long cap) {
  try {
    const FusedPlan& fp = cached_fused(p, min_threads);
    out[0] = fp.ok ? 1 : 0;
    out[1] = fp.anchor;
    out[2] = fp.t_lo;
    out[3] = fp.t_hi;
    out[4] = fp.vs;
    out[5] = fp.ks;
    out[6] = (long long) fp.insts.size();
    out[7] = fp.n_inst_home;
    out[8] = (long long) fp.fin.size();
    out[9] = 0;
    out[10] = 0;
    for (const auto& pc: fp.pieces) {
      out[9] += (long long) pc.fout.size();
      out[10] += pc.prog.n_fma;
    }
    out[11] = (long long) fp.pieces.size();
    if (why && cap > 0) {
      const size_t n = std::min<size_t>(fp.why.size(), (size_t) cap - 1);
      std::memcpy(why, fp.why.data(), n);
      why[n] = 0;
    }
    return AB_OK;
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}

int ab_problem_fused_layout(const ab_problem* p, int min_threads, long long out[4], int* ex_lo, int* ex_hi) {
  AB_TRY
  const FusedPlan& fp = cached_fused(p, min_threads);
  if (!fp.ok) return fail(AB_ERR_INVALID, "no fused interior: " + fp.why);
  out[0] = fp.vb;
  out[1] = fp.vs;
  out[2] = fp.kb;
  out[3] = fp.ks;
  for (size_t f = 0; f < p->p.funcs.size(); f++) {
    if (ex_lo) ex_lo[f] = fp.ex_lo[f];
    if (ex_hi) ex_hi[f] = fp.ex_hi[f];
  }
  return AB_OK;
  AB_CATCH
}
int ab_problem_dest_maps(const ab_problem* p, int f, int* kkt_dest, int* f_dest, long long out[2]) {
  AB_TRY
  const Problem& P = p->p;
  if (!P.analyzed) return fail(AB_ERR_INVALID, "ab_problem_analyze() first");
  const FuncRec& fr = P.funcs.at((size_t) f);
  const long long nt = (long long) fr.nappl * fr.triplets_per_app();
  if (kkt_dest) std::copy(P.kkt_dest.begin() + P.trip_base[(size_t) f], P.kkt_dest.begin() + P.trip_base[(size_t) f] + nt, kkt_dest);
  if (f_dest && fr.role != 0)
    std::copy(P.f_dest.begin() + P.fdest_base[(size_t) f], P.f_dest.begin() + P.fdest_base[(size_t) f] + (long long) fr.nappl * fr.orows, f_dest);
  out[0] = P.g_stage.at((size_t) f);
  out[1] = P.v_stage.at((size_t) f);
  return AB_OK;
  AB_CATCH
}
long long ab_problem_fused_residual_ptx(const ab_problem* p, int min_threads, int mode, char* buf, long long cap) {
  try {
    if (mode < 4 || mode > 7) return fail(AB_ERR_INVALID, "residual-only modes are AB_EVAL_RHS .. AB_EVAL_OBJ");
    const FusedPlan& fp = cached_fused(p, min_threads);
    if (!fp.ok || fp.residual[mode].empty()) return 0;
    const std::string s = fused_ptx(fp, 0, nullptr, nullptr, 0, mode);
    if (buf && cap > 0) {
      const size_t n = std::min<size_t>(s.size(), (size_t) cap - 1);
      std::memcpy(buf, s.data(), n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}
long long ab_problem_fused_mode_ptx(const ab_problem* p, int min_threads, int mode, int piece, char* buf, long long cap) {
  try {
    if (mode < 1 || mode > 7) return fail(AB_ERR_INVALID, "modes AB_EVAL_KKTNO .. AB_EVAL_OBJ (evalKKT: ab_problem_fused_ptx)");
    const FusedPlan& fp = cached_fused(p, min_threads);
    if (!fp.ok || piece < 0 || piece >= (int) fp.residual[mode].size()) return 0;
    const std::string s = fused_ptx(fp, piece, nullptr, nullptr, 0, mode);
    if (buf && cap > 0) {
      const size_t n = std::min<size_t>(s.size(), (size_t) cap - 1);
      std::memcpy(buf, s.data(), n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}
long long ab_problem_fused_ptx(const ab_problem* p, int min_threads, int piece, char* buf, long long cap) {
  try {
    const FusedPlan& fp = cached_fused(p, min_threads);
    if (!fp.ok || piece < 0 || piece >= (int) fp.pieces.size()) return 0;
    const std::string s = fused_ptx(fp, piece, nullptr, nullptr);
    if (buf && cap > 0) {
      const size_t n = std::min<size_t>(s.size(), (size_t) cap - 1);
      std::memcpy(buf, s.data(), n);
      buf[n] = 0;
    }
    return (long long) s.size();
  } catch (const std::exception& e) {
    return fail(AB_ERR_INVALID, e.what());
  }
}

// ------------------------------------------------------------------ plan
#define PLAN_CUDA(call)                                                              \
  do {                                                                               \
    cudaError_t e__ = (call);                                                        \
    if (e__ != cudaSuccess) return fail((int) e__, std::string(#call) + ": " + cudaGetErrorString(e__)); \
  } while (0)

int ab_plan_create(const ab_problem* p, int device, ab_plan** out) {
  AB_TRY
  if (!out) return fail(AB_ERR_INVALID, "null out pointer");
  *out = nullptr;
  if (ab_device_count() <= 0)
    return fail(AB_ERR_NO_DEVICE, "no CUDA device: the KKT evaluation path has no CPU fallback");
  std::unique_ptr<ab_plan> pl(new ab_plan());
  pl->plan.prob = p->p;
  int rc = pl->plan.build(device);
  if (rc != 0) return fail(rc, pl->plan.last_error);
  *out = pl.release();
  return AB_OK;
  AB_CATCH
}
int ab_plan_create_sharded(const ab_problem* p, int device, int rank, int world, ab_plan** out) {
  AB_TRY
  if (!out) return fail(AB_ERR_INVALID, "null out pointer");
  *out = nullptr;
  if (world < 1 || rank < 0 || rank >= world) return fail(AB_ERR_INVALID, "ab_plan_create_sharded: bad rank / world");
  if (ab_device_count() <= 0)
    return fail(AB_ERR_NO_DEVICE, "no CUDA device: the KKT evaluation path has no CPU fallback");
  std::unique_ptr<ab_plan> pl(new ab_plan());
  pl->plan.prob = p->p;
  int rc = pl->plan.build(device, rank, world);
  if (rc != 0) return fail(rc, pl->plan.last_error);
  *out = pl.release();
  return AB_OK;
  AB_CATCH
}
void ab_plan_destroy(ab_plan* plan) { delete plan; }
long long ab_plan_cut_count(const ab_plan* plan) { return plan->plan.active_cut_count(); }
int ab_plan_fused_segments(const ab_plan* plan) {
  if (plan->plan.windowed) return plan->plan.last_fused ? (int) (plan->plan.win_f_e - plan->plan.win_f_b) : 0;
  return plan->plan.last_fused ? plan->plan.fused.t_hi - plan->plan.fused.t_lo : 0;
}
int ab_plan_output_ranges(const ab_plan* plan, int array, long long* out, int cap) {
  if (array < 0 || array >= OA_COUNT) return fail(AB_ERR_INVALID, "bad array");
  const auto& R = plan->plan.out_ranges[array];
  for (size_t k = 0; k < R.size() && (int) k < cap; k++) {
    out[2 * k] = R[k][0];
    out[2 * k + 1] = R[k][1];
  }
  return (int) R.size();
}
int ab_plan_get_outputs_packed(ab_plan* plan, double* kkt, double* pgx, double* agx, double* fxe, double* fxi, double* val) {
  Plan& P = plan->plan;
  if (P.world <= 1) return fail(AB_ERR_INVALID, "ab_plan_get_outputs_packed: not a sharded plan");
  PLAN_CUDA(cudaSetDevice(P.device));
  double* dst[OA_COUNT] = {kkt, pgx, agx, fxe, fxi, val};
  const double* src[OA_COUNT] = {P.d_kkt, P.d_pgx, P.d_agx, P.d_fxe, P.d_fxi, P.d_val};
  for (int a = 0; a < OA_COUNT; a++) {
    if (!dst[a]) continue;
    size_t off = 0;
    for (const auto& r: P.out_ranges[a]) {
      const size_t n = (size_t) (r[1] - r[0]);
      PLAN_CUDA(cudaMemcpyAsync(dst[a] + off, src[a] + r[0], n * sizeof(double), cudaMemcpyDeviceToHost, P.stream));
      off += n;
    }
  }
  PLAN_CUDA(cudaStreamSynchronize(P.stream));
  return AB_OK;
}
void* ab_plan_cut_buffer(ab_plan* plan) { return plan->plan.d_cutbuf; }
int ab_plan_cut_pack(ab_plan* plan) {
  AB_TRY
  int rc = plan->plan.cut_pack();
  if (rc != 0) return fail(rc, plan->plan.last_error);
  return AB_OK;
  AB_CATCH
}
int ab_plan_cut_unpack(ab_plan* plan) {
  AB_TRY
  int rc = plan->plan.cut_unpack();
  if (rc != 0) return fail(rc, plan->plan.last_error);
  return AB_OK;
  AB_CATCH
}

// ------------------------------------------------------------------ partition (host only; also used by the CPU tests)
struct ab_partition {
  RankPartition part;
};
ab_partition* ab_partition_create(const ab_problem* p, int rank, int world) {
  try {
    std::unique_ptr<ab_partition> q(new ab_partition());
    q->part = make_partition(p->p, rank, world);
    return q.release();
  } catch (const std::exception& e) {
    fail(AB_ERR_INVALID, e.what());
    return nullptr;
  }
}
void ab_partition_destroy(ab_partition* q) { delete q; }
int ab_partition_app_range(const ab_partition* q, int f, int out[2]) {
  AB_TRY
  out[0] = q->part.app_begin.at(f);
  out[1] = q->part.app_end.at(f);
  return AB_OK;
  AB_CATCH
}
int ab_partition_gather_sizes(const ab_partition* q, int array, int level, long long out[2]) {
  AB_TRY
  if (array < 0 || array >= OA_COUNT || level < 0 || level > 1) return fail(AB_ERR_INVALID, "bad array / level");
  out[0] = q->part.g[array][level].ntargets();
  out[1] = (long long) q->part.g[array][level].src.size();
  return AB_OK;
  AB_CATCH
}
int ab_partition_gather(const ab_partition* q, int array, int level, long long* ptr, int* src, int* target) {
  AB_TRY
  if (array < 0 || array >= OA_COUNT || level < 0 || level > 1) return fail(AB_ERR_INVALID, "bad array / level");
  const GatherList& g = q->part.g[array][level];
  std::memcpy(ptr, g.ptr.data(), g.ptr.size() * sizeof(long long));
  if (!g.src.empty()) std::memcpy(src, g.src.data(), g.src.size() * sizeof(int));
  if (!g.target.empty()) std::memcpy(target, g.target.data(), g.target.size() * sizeof(int));
  return AB_OK;
  AB_CATCH
}
long long ab_partition_cut_size(const ab_partition* q, int array) {
  if (array < 0 || array >= OA_COUNT) return -1;
  return (long long) q->part.cut_target[array].size();
}
int ab_partition_cut(const ab_partition* q, int array, int* target, int* owner, int* mine) {
  AB_TRY
  if (array < 0 || array >= OA_COUNT) return fail(AB_ERR_INVALID, "bad array");
  const size_t n = q->part.cut_target[array].size();
  for (size_t k = 0; k < n; k++) {
    target[k] = q->part.cut_target[array][k];
    owner[k] = q->part.cut_owner[array][k];
    mine[k] = q->part.cut_mine[array][k];
  }
  return AB_OK;
  AB_CATCH
}
/* the problem's own (unpartitioned) gather lists and staging layout, for the tests' emulation of a sharded evaluation */
int ab_problem_gather_sizes(const ab_problem* p, int array, int level, long long out[3]) {
  AB_TRY
  if (array < 0 || array >= OA_COUNT || level < 0 || level > 1) return fail(AB_ERR_INVALID, "bad array / level");
  const GatherList* gl[OA_COUNT] = {p->p.g_kkt, p->p.g_pgx, p->p.g_agx, p->p.g_fxe, p->p.g_fxi, p->p.g_val};
  out[0] = gl[array][level].ntargets();
  out[1] = (long long) gl[array][level].src.size();
  out[2] = p->p.stage_size;
  return AB_OK;
  AB_CATCH
}
int ab_problem_gather(const ab_problem* p, int array, int level, long long* ptr, int* src, int* target) {
  AB_TRY
  if (array < 0 || array >= OA_COUNT || level < 0 || level > 1) return fail(AB_ERR_INVALID, "bad array / level");
  const GatherList* gl[OA_COUNT] = {p->p.g_kkt, p->p.g_pgx, p->p.g_agx, p->p.g_fxe, p->p.g_fxi, p->p.g_val};
  const GatherList& g = gl[array][level];
  std::memcpy(ptr, g.ptr.data(), g.ptr.size() * sizeof(long long));
  if (!g.src.empty()) std::memcpy(src, g.src.data(), g.src.size() * sizeof(int));
  if (!g.target.empty()) std::memcpy(target, g.target.data(), g.target.size() * sizeof(int));
  return AB_OK;
  AB_CATCH
}

int ab_plan_set_inputs(ab_plan* plan, const double* X, const double* LE, const double* LI) {
  Plan& P = plan->plan;
  PLAN_CUDA(cudaSetDevice(P.device));
  if (X) PLAN_CUDA(cudaMemcpyAsync(P.d_X, X, sizeof(double) * P.prob.nvars, cudaMemcpyHostToDevice, P.stream));
  if (LE && P.prob.neq) PLAN_CUDA(cudaMemcpyAsync(P.d_LE, LE, sizeof(double) * P.prob.neq, cudaMemcpyHostToDevice, P.stream));
  if (LI && P.prob.niq) PLAN_CUDA(cudaMemcpyAsync(P.d_LI, LI, sizeof(double) * P.prob.niq, cudaMemcpyHostToDevice, P.stream));
  return AB_OK;
}
long long ab_plan_input_ranges(const ab_plan* plan, int which, long long* out, long long cap) {
  if (which < 0 || which > 2) return fail(AB_ERR_INVALID, "which must be 0 (X), 1 (LE) or 2 (LI)");
  const auto& r = plan->plan.in_ranges[which];
  for (size_t k = 0; k < r.size() && (long long) k < cap; k++) {
    out[2 * k] = r[k].first;
    out[2 * k + 1] = r[k].second;
  }
  return (long long) r.size();
}
int ab_plan_set_inputs_sliced(ab_plan* plan, const double* X, const double* LE, const double* LI) {
  Plan& P = plan->plan;
  PLAN_CUDA(cudaSetDevice(P.device));
  const double* host[3] = {X, LE, LI};
  double* dev[3] = {P.d_X, P.d_LE, P.d_LI};
  for (int w = 0; w < 3; w++) {
    if (!host[w]) continue;
    for (const auto& r: P.in_ranges[w])
      PLAN_CUDA(cudaMemcpyAsync(dev[w] + r.first, host[w] + r.first, sizeof(double) * (size_t) (r.second - r.first), cudaMemcpyHostToDevice,
                                P.stream));
  }
  return AB_OK;
}
int ab_plan_set_solver_coeffs(ab_plan* plan, const double* c) {
  int rc = plan->plan.set_solver_coeffs(c);
  return rc ? fail(rc, plan->plan.last_error) : AB_OK;
}
int ab_plan_eval(ab_plan* plan, int mode, double obj_scale) {
  int rc = plan->plan.eval(mode, obj_scale);
  return rc ? fail(rc, plan->plan.last_error) : AB_OK;
}
int ab_plan_sync(ab_plan* plan) {
  PLAN_CUDA(cudaSetDevice(plan->plan.device));
  PLAN_CUDA(cudaStreamSynchronize(plan->plan.stream));
  return AB_OK;
}
int ab_plan_get_outputs(ab_plan* plan, double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt) {
  Plan& P = plan->plan;
  PLAN_CUDA(cudaSetDevice(P.device));
  auto D2H = [&](double* dst, const double* src, size_t n) -> cudaError_t {
    if (!dst || n == 0) return cudaSuccess;
    return cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToHost, P.stream);
  };
  if (P.world > 1) {
    // sharded plan: only the ranges this rank can write travel (everything else is zero here by construction and is
    // left untouched in the caller's arrays)
    double* dst[OA_COUNT] = {kkt, pgx, agx, fxe, fxi, val};
    const double* src[OA_COUNT] = {P.d_kkt, P.d_pgx, P.d_agx, P.d_fxe, P.d_fxi, P.d_val};
    for (int a = 0; a < OA_COUNT; a++)
      if (dst[a])
        for (const auto& r: P.out_ranges[a]) PLAN_CUDA(D2H(dst[a] + r[0], src[a] + r[0], (size_t) (r[1] - r[0])));
    PLAN_CUDA(cudaStreamSynchronize(P.stream));
    return AB_OK;
  }
  PLAN_CUDA(D2H(val, P.d_val, 1));
  PLAN_CUDA(D2H(pgx, P.d_pgx, P.prob.nvars));
  PLAN_CUDA(D2H(agx, P.d_agx, P.prob.nvars));
  PLAN_CUDA(D2H(fxe, P.d_fxe, P.prob.neq));
  PLAN_CUDA(D2H(fxi, P.d_fxi, P.prob.niq));
  PLAN_CUDA(D2H(kkt, P.d_kkt, (size_t) P.prob.nnz()));
  PLAN_CUDA(cudaStreamSynchronize(P.stream));
  return AB_OK;
}
int ab_plan_eval_host(ab_plan* plan, int mode, double obj_scale, const double* X, const double* LE, const double* LI, double* val,
                      double* pgx, double* agx, double* fxe, double* fxi, double* kkt) {
  int rc = ab_plan_set_inputs(plan, X, LE, LI);
  if (rc) return rc;
  rc = ab_plan_eval(plan, mode, obj_scale);
  if (rc) return rc;
  return ab_plan_get_outputs(plan, val, pgx, agx, fxe, fxi, kkt);
}
int ab_plan_check_guards(ab_plan* plan) {
  const int rc = plan->plan.check_guards();
  if (rc != 0) fail(rc, plan->plan.last_error);
  return rc;
}
void* ab_plan_device_ptr(ab_plan* plan, int which) {
  Plan& P = plan->plan;
  switch (which) {
    case AB_PTR_X: return P.d_X;
    case AB_PTR_LE: return P.d_LE;
    case AB_PTR_LI: return P.d_LI;
    case AB_PTR_KKT: return P.d_kkt;
    case AB_PTR_FXE: return P.d_fxe;
    case AB_PTR_FXI: return P.d_fxi;
    case AB_PTR_PGX: return P.d_pgx;
    case AB_PTR_AGX: return P.d_agx;
    case AB_PTR_VAL: return P.d_val;
    default: return nullptr;
  }
}
int ab_plan_profile(ab_plan* plan, int enable) {
  plan->plan.profiling = enable;
  plan->plan.prof_labels.clear();
  return AB_OK;
}
int ab_plan_profile_read(ab_plan* plan, float* ms, int cap) {
  int n = plan->plan.profile_read(ms, cap);
  return n;
}
long long ab_plan_profile_label(ab_plan* plan, int k, char* buf, long long cap) {
  if (k < 0 || k >= (int) plan->plan.prof_labels.size()) return fail(AB_ERR_INVALID, "launch index out of range");
  const std::string& s = plan->plan.prof_labels[k];
  if (buf && cap > 0) {
    long long n = std::min<long long>(cap - 1, (long long) s.size());
    std::memcpy(buf, s.data(), (size_t) n);
    buf[n] = 0;
  }
  return (long long) s.size();
}
void* ab_plan_stream(ab_plan* plan) { return (void*) plan->plan.stream; }
long long ab_plan_launches(const ab_plan* plan) { return plan->plan.launches; }
long long ab_plan_specialised_launches(const ab_plan* plan) { return plan->plan.jit_launches; }
int ab_plan_kernel_info(const ab_plan* plan, int k, int mode, int out[4]) {
  AB_TRY
  const Plan& P = plan->plan;
  if (mode < 0 || mode > 4) throw std::invalid_argument("mode must be 0..4");
  out[0] = P.jit_state[0].at(k)[mode];
  out[1] = P.jits[0].at(k)[mode].nregs;
  out[2] = P.jits[0].at(k)[mode].spill_bytes;
  out[3] = P.affs.at(k).a1 - P.affs.at(k).a0;
  return AB_OK;
  AB_CATCH
}

}  // extern "C"
