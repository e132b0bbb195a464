#include "plan.hpp"

#include <algorithm>
#include <stdexcept>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

namespace ab {

#define AB_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) {                                                \
      last_error = std::string(#call) + ": " + cudaGetErrorString(e__);      \
      return (int) e__;                                                      \
    }                                                                        \
  } while (0)

// ASSET_B200_GUARD=1 (checking aid; compute-sanitizer is not available on every pool): every device allocation of the
// plan is wrapped in two guard bands of GUARD_BYTES filled with a byte pattern, and Plan::check_guards() reports the
// first band a kernel wrote into.  Together with the comparison of every output with the oracle this covers what
// memcheck would: a store outside an array lands in a band (or in another array's interior, where the oracle sees it).
static constexpr size_t GUARD_BYTES = 1 << 16;
static constexpr unsigned char GUARD_PATTERN = 0xA5;
template<class T>
T* Plan::dalloc(size_t n) {
  static const bool guard = getenv("ASSET_B200_GUARD") != nullptr;
  void* p = nullptr;
  if (n == 0) n = 1;
  if (!guard) {
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
    owned.push_back(p);
    return (T*) p;
  }
  const size_t bytes = (n * sizeof(T) + 255) / 256 * 256;
  if (cudaMalloc(&p, bytes + 2 * GUARD_BYTES) != cudaSuccess) return nullptr;
  owned.push_back(p);
  cudaMemset(p, GUARD_PATTERN, GUARD_BYTES);
  cudaMemset((char*) p + GUARD_BYTES + n * sizeof(T), GUARD_PATTERN, bytes - n * sizeof(T) + GUARD_BYTES);
  guards.push_back({(char*) p, (char*) p + GUARD_BYTES + n * sizeof(T), bytes - n * sizeof(T) + GUARD_BYTES});
  return (T*) ((char*) p + GUARD_BYTES);
}

// 0: every guard band is intact; k > 0: the k-th allocation's band was written (message in last_error); < 0 CUDA error
int Plan::check_guards() {
  if (guards.empty()) {
    last_error = "guard bands are off (set ASSET_B200_GUARD=1 before the plan is created)";
    return -1;
  }
  cudaDeviceSynchronize();
  std::vector<unsigned char> h;
  for (size_t k = 0; k < guards.size(); k++) {
    const Guard& g = guards[k];
    for (int side = 0; side < 2; side++) {
      const char* at = side ? g.hi : g.lo;
      const size_t nb = side ? g.hi_bytes : GUARD_BYTES;
      h.resize(nb);
      if (cudaMemcpy(h.data(), at, nb, cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
      for (size_t b = 0; b < nb; b++)
        if (h[b] != GUARD_PATTERN) {
          last_error = "allocation " + std::to_string(k) + ": the guard band " + (side ? "after" : "before") + " it was written at byte offset "
                       + std::to_string(side ? (long long) b : (long long) b - (long long) GUARD_BYTES);
          return (int) k + 1;
        }
    }
  }
  return 0;
}
template<class T>
T* Plan::dupload(const std::vector<T>& v) {
  T* p = dalloc<T>(v.size());
  if (p && !v.empty()) {
    if (cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) return nullptr;
  }
  return p;
}

Plan::~Plan() {
  for (auto& v: jits)
    for (auto& a: v)
      for (JitKernel& k: a) unload_jit(k);
  for (JitKernel& k: fused_jits) unload_jit(k);
  for (auto& v: fused_res_jits)
    for (JitKernel& k: v) unload_jit(k);
  for (EvalGraph& g: graphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  if (ev_fork) cudaEventDestroy(ev_fork);
  if (ev_join) cudaEventDestroy(ev_join);
  for (int i = 1; i < SIDE_POOL; i++)
    if (side_pool[i]) cudaStreamDestroy(side_pool[i]);
  for (int i = 0; i < SIDE_POOL; i++)
    if (ev_pool[i]) cudaEventDestroy(ev_pool[i]);
  if (ev_ends) cudaEventDestroy(ev_ends);
  if (side) cudaStreamDestroy(side);
  if (bulk2) cudaStreamDestroy(bulk2);
  if (ev_join2) cudaEventDestroy(ev_join2);
  for (cudaEvent_t e: prof_events) cudaEventDestroy(e);
  for (void* p: owned) cudaFree(p);
  if (stream) cudaStreamDestroy(stream);
}

// Largest application range around the middle of the function over which every index map is affine in the application
// number (true for all interior mesh segments of a phase: the transcription and the CSR pattern are periodic there).
static AffineMaps find_affine(const int* V, int ir, const int* C, int orr, const int* FD, const int* KD, int E, int napp) {
  AffineMaps am;
  if (napp < 8) return am;
  const long long mid = napp / 2;
  struct Col {
    const int* p;
    int w;
    std::vector<int>*base, *stride;
  };
  std::vector<Col> cols;
  cols.push_back({V, ir, &am.vbase, &am.vstride});
  if (C) cols.push_back({C, orr, &am.cbase, &am.cstride});
  if (FD) cols.push_back({FD, orr, &am.fbase, &am.fstride});
  cols.push_back({KD, E, &am.kbase, &am.kstride});
  for (Col& c: cols) {
    c.base->resize(c.w);
    c.stride->resize(c.w);
    for (int q = 0; q < c.w; q++) {
      const long long st = (long long) c.p[(mid + 1) * c.w + q] - c.p[mid * c.w + q];
      if (st >= (1 << 23) || st <= -(1 << 23)) return AffineMaps();
      (*c.stride)[q] = (int) st;
    }
  }
  auto row_ok = [&](long long a) {
    for (const Col& c: cols)
      for (int q = 0; q < c.w; q++)
        if ((long long) c.p[a * c.w + q] != (long long) c.p[mid * c.w + q] + (a - mid) * (*c.stride)[q]) return false;
    return true;
  };
  long long a0 = mid, a1 = mid + 1;
  while (a0 > 0 && row_ok(a0 - 1)) a0--;
  while (a1 < napp && row_ok(a1)) a1++;
  if (a1 - a0 < 4) return AffineMaps();
  for (Col& c: cols)
    for (int q = 0; q < c.w; q++) (*c.base)[q] = c.p[a0 * c.w + q];
  am.a0 = (int) a0;
  am.a1 = (int) a1;
  return am;
}

AffineMaps function_affine(const Problem& prob, int fi) {
  const FuncRec& fr = prob.funcs.at(fi);
  if (fr.nappl < 8) return AffineMaps();
  return find_affine(fr.vindex.data(), fr.ir, fr.role ? fr.cindex.data() : nullptr, fr.orows,
                     fr.role ? prob.f_dest.data() + prob.fdest_base[fi] : nullptr, prob.kkt_dest.data() + prob.trip_base[fi],
                     fr.triplets_per_app(), fr.nappl);
}

// Outputs of a program that are the constant zero: (output kind, index) pairs.
static std::vector<std::pair<uint32_t, uint32_t>> const_zero_outputs(const Program& P) {
  std::vector<std::pair<uint32_t, uint32_t>> z;
  const size_t nb = P.code.size() / 4;
  for (size_t bi = 0; bi < nb; bi++) {
    const PInstr* ins = &P.code[4 * bi];
    const uint32_t kind = ins[0].d >> 28, cnt = (ins[0].d >> 24) & 15u;
    if (kind == PK_END) break;
    if (kind != PK_OUT) continue;
    for (uint32_t q = 0; q < cnt; q++)
      if (!(ins[q].a & P_SLOT) && P.consts.at(ins[q].a >> 3) == 0.0) z.emplace_back(ins[q].b, ins[q].c);
  }
  return z;
}

// Drops the sources flagged in `dead` (staging slots that only ever hold a structural zero) from a gather list.
static GatherList prune_gather(const GatherList& g, const std::vector<unsigned char>& dead) {
  GatherList o;
  o.ptr.assign(1, 0);
  o.target = g.target;
  o.src.reserve(g.src.size());
  for (int t = 0; t < g.ntargets(); t++) {
    for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1]; k++)
      if (!dead[(size_t) g.src[(size_t) k]]) o.src.push_back(g.src[(size_t) k]);
    o.ptr.push_back((long long) o.src.size());
  }
  return o;
}

static bool up_gather(Plan& P, const GatherList gl[2], DevGather& dg) {
  for (int w = 0; w < 2; w++) {
    dg.dev[w].ntargets = gl[w].ntargets();
    long long maxfan = 0;
    for (int t = 0; t < gl[w].ntargets(); t++) maxfan = std::max(maxfan, gl[w].ptr[(size_t) t + 1] - gl[w].ptr[t]);
    dg.dev[w].maxparts = (int) std::max<long long>(1, (maxfan + GATHER_PART - 1) / GATHER_PART);
    long long* ptr = P.dupload(gl[w].ptr);
    int* src = P.dupload(gl[w].src);
    int* tgt = P.dupload(gl[w].target);
    if (!ptr || !src || !tgt) return false;
    dg.dev[w].ptr = ptr;
    dg.dev[w].src = src;
    dg.dev[w].target = tgt;
  }
  return true;
}

// Windowed plan: of the fused evaluation's gather lists keep what this rank owns, and make the cut set.
//   * a target in a row of a mesh variable or of a constraint the rank owns stays as it is -- every source of it must
//     be something the rank evaluates (checked);
//   * a target in a row this rank does not own (the window's halo) goes;
//   * a *far* target -- rows of parameter variables, accumulated constraint rows, the objective value -- stays on every
//     rank with the sources the rank itself writes (the solver coefficients on rank 0 only) and joins the cut set, which
//     is summed over the ranks after the evaluation; far targets are the same on every rank, in the same order, because
//     they do not depend on the window.
// Also fills own_app_lo / own_app_hi, the fused thread range, and the output blocks.
void Plan::window_filter(GatherList (*fl)[2], GatherList (*flE)[2], std::vector<int>* cut_t, int true_world) {
  const Problem::Window& W = prob.win;
  const long long R = fused.t_hi - fused.t_lo;
  const long long abase = W.app_base[(size_t) fused.anchor];
  // fused threads of owned segments, in thread numbers relative to t_lo
  win_f_b = std::min(std::max<long long>(W.seg_lo - abase - fused.t_lo, 0), R);
  win_f_e = std::min(std::max<long long>(W.seg_hi - abase - fused.t_lo, 0), R);
  if (!W.first && W.seg_lo - abase < fused.t_lo) throw std::runtime_error("window: the halo is shorter than the non-periodic margin of the mesh");
  if (!W.last && W.seg_hi - abase > fused.t_hi) throw std::runtime_error("window: the halo is shorter than the non-periodic margin of the mesh");
  // staging slots this rank writes in an evaluation
  std::vector<unsigned char> wr((size_t) prob.stage_size + 1, 0);
  std::vector<unsigned char> own_eq((size_t) std::max(prob.neq, 1), 0), own_iq((size_t) std::max(prob.niq, 1), 0);
  for (size_t k = 0; k < fns.size(); k++) {
    const int fi = prob.order[k];
    const FuncRec& fr = prob.funcs[fi];
    const int E = fr.triplets_per_app();
    for (int ap = W.own_lo[(size_t) fi]; ap < W.own_hi[(size_t) fi]; ap++) {
      if (fr.role != 0 && fr.unique_constraints)
        for (int o = 0; o < fr.orows; o++) (fr.role == 1 ? own_eq : own_iq)[(size_t) fr.cindex[(size_t) ap * fr.orows + o]] = 1;
      if (ap >= fused.ex_lo[fi] && ap < fused.ex_hi[fi]) continue;  // a fused thread evaluates it
      for (int e = 0; e < E; e++) {
        const int d = prob.kkt_dest[(size_t) (prob.trip_base[fi] + (long long) ap * E + e)];
        if (d < 0) wr[(size_t) ~d] = 1;
      }
      if (fr.role != 0)
        for (int o = 0; o < fr.orows; o++) {
          const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) ap * fr.orows + o)];
          if (d < 0) wr[(size_t) ~d] = 1;
        }
      for (int i = 0; i < fr.ir; i++) wr[(size_t) (prob.g_stage[fi] + (long long) ap * fr.ir + i)] = 1;
      if (fr.role == 0) wr[(size_t) (prob.v_stage[fi] + ap)] = 1;
    }
  }
  for (const auto& pc: fused.pieces)
    for (const JitOut& w: pc.fout)
      if (w.arr == FA_STAGE)
        for (long long r = win_f_b; r < win_f_e; r++) wr[(size_t) (w.base + r * w.stride)] = 1;
  const long long s0 = prob.stage_solver, s1 = prob.stage_solver + prob.n_solver;
  const int PV = prob.nvars, SV = prob.niq, EC = prob.neq;
  // constraint rows that several applications add into (accumulations) belong to no rank
  std::vector<unsigned char> acc_eq((size_t) std::max(prob.neq, 1), 0), acc_iq((size_t) std::max(prob.niq, 1), 0);
  for (const FuncRec& fr: prob.funcs)
    if (fr.role != 0 && !fr.unique_constraints)
      for (int c: fr.cindex) (fr.role == 1 ? acc_eq : acc_iq)[(size_t) c] = 1;
  // 0 not owned, 1 owned, 2 far (every rank contributes, summed through the cut buffer)
  auto var_class = [&](long long v) { return W.is_far(v) ? 2 : (W.owns_var(v) ? 1 : 0); };
  auto con_class = [&](long long r) {  // row / column index >= nvars: slack, equality or inequality row
    const bool iq = r < PV + SV || r >= PV + SV + EC;
    const long long c = r < PV + SV ? r - PV : (r < PV + SV + EC ? r - PV - SV : r - PV - SV - EC);
    if ((iq ? acc_iq : acc_eq)[(size_t) c]) return 2;
    return (iq ? own_iq : own_eq)[(size_t) c] ? 1 : 0;
  };
  // KKT entry (r, c), r <= c: the row of an ordinary mesh variable belongs to its segment's rank; in the row of a far
  // variable an entry goes with its column -- another far variable: far; an ordinary variable or a constraint: that one's owner
  auto entry_class = [&](long long r, long long c) {
    if (r >= PV) return con_class(r);
    if (!W.is_far(r)) return W.owns_var(r) ? 1 : 0;
    if (c < PV) return var_class(c);
    return con_class(c);
  };
  for (int a = 0; a < OA_COUNT; a++) {
    cut_t[a].clear();
    for (int which = 0; which < 2; which++)
      for (int w = 0; w < 2; w++) {
        GatherList& g = which ? flE[a][w] : fl[a][w];
        GatherList o;
        o.ptr.assign(1, 0);
        for (int t = 0; t < g.ntargets(); t++) {
          const long long tg = g.target[(size_t) t];
          int cls;
          if (a == OA_KKT)
            cls = entry_class((long long) (std::upper_bound(prob.indptr.begin(), prob.indptr.end(), (int) tg) - prob.indptr.begin()) - 1,
                              prob.indices[(size_t) tg]);
          else if (a == OA_PGX || a == OA_AGX) cls = var_class(tg);
          else if (a == OA_FXE || a == OA_FXI)  // accumulated rows are far; rows nobody in the window writes are not this rank's
            cls = (a == OA_FXE ? acc_eq : acc_iq)[(size_t) tg] ? 2 : 0;
          else cls = 2;  // the objective value
          if (cls == 0) continue;
          for (long long q = g.ptr[(size_t) t]; q < g.ptr[(size_t) t + 1]; q++) {
            const int sidx = g.src[(size_t) q];
            const bool solver = sidx >= s0 && sidx < s1;
            if (cls == 1) {
              if (!solver && !wr[(size_t) sidx]) throw std::runtime_error("window: an owned output has a source this rank does not evaluate (halo too short)");
              o.src.push_back(sidx);
            } else if (solver ? rank == 0 : wr[(size_t) sidx] != 0) {
              o.src.push_back(sidx);
            }
          }
          o.ptr.push_back((long long) o.src.size());
          o.target.push_back((int) tg);
          if (cls == 2) cut_t[a].push_back((int) tg);
        }
        g = std::move(o);
      }
    // far targets must sit in the late lists (they are summed after the fused kernels) and be identical on every rank
    std::sort(cut_t[a].begin(), cut_t[a].end());
  }
  (void) true_world;
  // what this rank writes: output blocks
  const long long sizes[OA_COUNT] = {prob.nnz(), prob.nvars, prob.nvars, prob.neq, prob.niq, 1};
  for (int a = 0; a < OA_COUNT; a++) out_blocks[a].assign((size_t) ((sizes[a] + OUT_BLOCK - 1) / OUT_BLOCK) + 1, 0);
  auto touch = [&](int a, long long i) { out_blocks[a][(size_t) (i / OUT_BLOCK)] = 1; };
  auto arr_of = [](int fa) {
    return fa == FA_KKT ? (int) OA_KKT : fa == FA_FXE ? (int) OA_FXE : fa == FA_FXI ? (int) OA_FXI
           : fa == FA_PGX ? (int) OA_PGX : fa == FA_AGX ? (int) OA_AGX : -1;
  };
  for (const auto& pc: fused.pieces)
    for (const JitOut& w: pc.fout) {
      const int a = arr_of(w.arr);
      if (a < 0 || win_f_e <= win_f_b) continue;
      const long long i0 = w.base + win_f_b * w.stride, i1 = w.base + (win_f_e - 1) * w.stride;
      for (long long i = std::min(i0, i1) / OUT_BLOCK; i <= std::max(i0, i1) / OUT_BLOCK; i++) out_blocks[a][(size_t) i] = 1;
    }
  for (int a = 0; a < OA_COUNT; a++)
    for (int w = 0; w < 2; w++) {
      for (int t: fl[a][w].target) touch(a, t);
      for (int t: flE[a][w].target) touch(a, t);
    }
  for (size_t k = 0; k < fns.size(); k++) {
    const int fi = prob.order[k];
    const FuncRec& fr = prob.funcs[fi];
    const int E = fr.triplets_per_app();
    for (int ap = W.own_lo[(size_t) fi]; ap < W.own_hi[(size_t) fi]; ap++) {
      if (ap >= fused.ex_lo[fi] && ap < fused.ex_hi[fi]) continue;
      for (int e = 0; e < E; e++) {
        const int d = prob.kkt_dest[(size_t) (prob.trip_base[fi] + (long long) ap * E + e)];
        if (d >= 0) touch(OA_KKT, d);
      }
      if (fr.role != 0)
        for (int o = 0; o < fr.orows; o++) {
          const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) ap * fr.orows + o)];
          if (d >= 0) touch(fr.role == 1 ? OA_FXE : OA_FXI, d);
        }
    }
  }
}

void Plan::release_host_analysis() {
  prob.nnz_kept = prob.nnz();
  auto drop = [](auto& v) {
    typename std::remove_reference<decltype(v)>::type().swap(v);
  };
  drop(prob.indices);
  drop(prob.kktloc);
  drop(prob.kkt_dest);
  drop(prob.f_dest);
  for (GatherList* g: {prob.g_kkt, prob.g_pgx, prob.g_agx, prob.g_fxe, prob.g_fxi, prob.g_val})
    for (int w = 0; w < 2; w++) {
      drop(g[w].ptr);
      drop(g[w].src);
      drop(g[w].target);
    }
  for (FuncRec& fr: prob.funcs) {
    drop(fr.vindex);
    drop(fr.cindex);
    drop(fr.appconst);
  }
}

int Plan::build(int dev, int rank_, int world_) {
  if (!prob.analyzed) prob.analyze();
  device = dev;
  rank = rank_;
  world = world_;
  // A windowed problem (Problem::window) is this rank's share already: the plan is built like a single-device plan on it,
  // then restricted to what the rank owns (window_filter below); rank / world matter for the cut set only.
  windowed = prob.win.on;
  const int true_world = world;
  if (windowed) world = 1;
  if (world > 1) {
    try {
      part = make_partition(prob, rank, world);
    } catch (const std::exception& e) {
      last_error = e.what();
      return (int) cudaErrorInvalidValue;
    }
  }
  AB_CUDA(cudaSetDevice(dev));
  AB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  int* d_kkt_dest = dupload(prob.kkt_dest);
  int* d_f_dest = dupload(prob.f_dest);
  if (!d_kkt_dest || !d_f_dest) {
    last_error = "out of device memory for the destination maps";
    return (int) cudaErrorMemoryAllocation;
  }
  fns.clear();
  progs.clear();
  affs.clear();
  host_progs.clear();
  for (int v = 0; v < 2; v++) {
    jits[v].clear();
    jit_state[v].clear();
  }
  {
    // the side stream carries latency-bound single-block launches (mesh ends, edge segments): highest priority, so their
    // blocks are placed as soon as a slot frees up beside the bulk kernels
    int prio_lo = 0, prio_hi = 0;
    AB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    AB_CUDA(cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, getenv("ASSET_B200_NO_PRIORITY") ? prio_lo : prio_hi));
  }
  {
    int prio_lo = 0, prio_hi = 0;
    AB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    side_pool[0] = side;
    for (int i = 1; i < SIDE_POOL; i++)
      AB_CUDA(cudaStreamCreateWithPriority(&side_pool[i], cudaStreamNonBlocking, getenv("ASSET_B200_NO_PRIORITY") ? prio_lo : prio_hi));
    for (int i = 0; i < SIDE_POOL; i++) AB_CUDA(cudaEventCreateWithFlags(&ev_pool[i], cudaEventDisableTiming));
    AB_CUDA(cudaEventCreateWithFlags(&ev_ends, cudaEventDisableTiming));
  }
  AB_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  AB_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  AB_CUDA(cudaStreamCreateWithFlags(&bulk2, cudaStreamNonBlocking));
  AB_CUDA(cudaEventCreateWithFlags(&ev_join2, cudaEventDisableTiming));
  if (const char* e = getenv("ASSET_B200_JIT")) use_jit = std::atoi(e) != 0;
  if (const char* e = getenv("ASSET_B200_JIT_MIN")) jit_min_apps = std::max(1, std::atoi(e));
  for (int fi: prob.order) {
    const FuncRec& fr = prob.funcs[fi];
    const Tape& t = prob.tapes.at(fr.tape);
    if (t.nin != fr.ir || t.nout != fr.orows || !t.fn) {
      last_error = "tape/function size mismatch in " + fr.name;
      return (int) cudaErrorInvalidValue;
    }
    FnDev f;
    std::memset(&f, 0, sizeof(f));
    // device programs: one per evaluation mode this role can be asked for
    std::array<ProgDev, 5> pd;
    std::memset(pd.data(), 0, sizeof(ProgDev) * 5);
    std::array<Program, 5> hp;
    const AffineMaps am = function_affine(prob, fi);
    try {
      ScalarFn work = *t.fn;
      const double* d_tables = nullptr;  // one copy of the function's interpolation tables for all its modes
      for (int mode = 0; mode < 5; mode++) {
        if (fr.role == ROLE_OBJ && (mode == FM_JAC || mode == FM_JAC_GRAD)) continue;
        ProgramSpec sp;
        sp.is_objective = fr.role == ROLE_OBJ;
        sp.linear = fr.linear;
        sp.want_g = (mode == FM_GRAD || mode == FM_JAC_GRAD || mode == FM_ALL);
        sp.want_j = (mode >= FM_JAC);
        sp.want_h = (mode == FM_ALL);
        Program P = emit_program(work, sp);
        LinkedProgram L = link_program(P, 220u * 1024u, &am);
        std::vector<uint4> code(L.code.size());
        std::memcpy(code.data(), L.code.data(), L.code.size() * sizeof(PInstr));
        pd[mode].code = dupload(code);
        pd[mode].consts = dupload(P.consts);
        if (!pd[mode].code || !pd[mode].consts) {
          last_error = "out of device memory for the program of " + fr.name;
          return (int) cudaErrorMemoryAllocation;
        }
        if (!P.tables.empty()) {
          if (!d_tables) d_tables = dupload(P.tables);
          if (!d_tables) {
            last_error = "out of device memory for the interpolation tables of " + fr.name;
            return (int) cudaErrorMemoryAllocation;
          }
          pd[mode].tables = d_tables;
        }
        pd[mode].nconst = (int) P.consts.size();
        pd[mode].nslots = L.nslots;
        pd[mode].nthreads = L.nthreads;
        pd[mode].nbundles = P.bundles();
        pd[mode].smem_bytes = L.smem_bytes;
        hp[mode] = std::move(P);
      }
    } catch (const std::exception& e) {
      last_error = std::string(e.what()) + " (" + fr.name + ")";
      return (int) cudaErrorInvalidValue;
    }
    progs.push_back(pd);
    host_progs.push_back(std::move(hp));
    affs.push_back(am);
    for (int v = 0; v < 2; v++) {
      jits[v].emplace_back();
      jit_state[v].push_back({0, 0, 0, 0, 0});
    }
    f.aff_begin = am.a0;
    f.aff_end = am.a1;
    f.ir = fr.ir;
    f.orows = fr.orows;
    f.nappconst = fr.nappconst;
    f.role = fr.role;
    f.linear = fr.linear ? 1 : 0;
    f.app_begin = world > 1 ? part.app_begin[fi] : 0;
    f.app_end = world > 1 ? part.app_end[fi] : fr.nappl;
    f.E = fr.triplets_per_app();
    f.vindex = dupload(fr.vindex);
    f.cindex = dupload(fr.cindex);
    f.appconst = dupload(fr.appconst);
    if (!f.vindex || !f.cindex || !f.appconst) {
      last_error = "out of device memory for index maps of " + fr.name;
      return (int) cudaErrorMemoryAllocation;
    }
    f.kkt_dest = d_kkt_dest + prob.trip_base[fi];
    f.f_dest = d_f_dest + prob.fdest_base[fi];
    f.g_stage = prob.g_stage[fi];
    f.v_stage = prob.v_stage[fi];
    fns.push_back(f);
  }
  {
    long long best = -1;
    heavy_fn = -1;
    for (size_t k = 0; k < fns.size(); k++) {
      const long long w = (long long) (fns[k].app_end - fns[k].app_begin) * (fns[k].E + fns[k].ir + fns[k].orows);
      if (w > best) {
        best = w;
        heavy_fn = (int) k;
      }
    }
    if (const char* e = getenv("ASSET_B200_OVERLAP"))
      if (std::atoi(e) == 0) heavy_fn = -1;
  }
  d_X = dalloc<double>(prob.nvars);
  d_LE = dalloc<double>(prob.neq);
  d_LI = dalloc<double>(prob.niq);
  d_kkt = dalloc<double>((size_t) prob.nnz());
  d_fxe = dalloc<double>(prob.neq);
  d_fxi = dalloc<double>(prob.niq);
  d_pgx = dalloc<double>(prob.nvars);
  d_agx = dalloc<double>(prob.nvars);
  d_val = dalloc<double>(1);
  d_stage = dalloc<double>((size_t) prob.stage_size + 1);  // + the trash slot of the fused evaluation's redirected maps
  if (!d_X || !d_LE || !d_LI || !d_kkt || !d_fxe || !d_fxi || !d_pgx || !d_agx || !d_val || !d_stage) {
    last_error = "out of device memory for the value arrays";
    return (int) cudaErrorMemoryAllocation;
  }
  AB_CUDA(cudaMemset(d_stage, 0, sizeof(double) * (size_t) std::max<long long>(prob.stage_size, 1)));
  AB_CUDA(cudaMemset(d_kkt, 0, sizeof(double) * (size_t) std::max<long long>(prob.nnz(), 1)));
  // outputs another rank produces read as zero here, so that the sum over ranks is the full result
  AB_CUDA(cudaMemset(d_fxe, 0, sizeof(double) * (size_t) std::max(prob.neq, 1)));
  AB_CUDA(cudaMemset(d_fxi, 0, sizeof(double) * (size_t) std::max(prob.niq, 1)));
  AB_CUDA(cudaMemset(d_pgx, 0, sizeof(double) * (size_t) std::max(prob.nvars, 1)));
  AB_CUDA(cudaMemset(d_agx, 0, sizeof(double) * (size_t) std::max(prob.nvars, 1)));
  AB_CUDA(cudaMemset(d_val, 0, sizeof(double)));
  const GatherList* gsrc[OA_COUNT] = {prob.g_kkt, prob.g_pgx, prob.g_agx, prob.g_fxe, prob.g_fxi, prob.g_val};
  if (world > 1) {
    for (int a = 0; a < OA_COUNT; a++) gsrc[a] = part.g[a];
    cut_count = part.cut_total();
    {
      const long long sizes[OA_COUNT] = {prob.nnz(), prob.nvars, prob.nvars, prob.neq, prob.niq, 1};
      for (int a = 0; a < OA_COUNT; a++) out_blocks[a].assign((size_t) ((sizes[a] + OUT_BLOCK - 1) / OUT_BLOCK) + 1, 0);
      auto touch = [&](int a, long long i) { out_blocks[a][(size_t) (i / OUT_BLOCK)] = 1; };
      for (int fi: prob.order) {
        const FuncRec& fr = prob.funcs[fi];
        const int E = fr.triplets_per_app();
        for (int ap = part.app_begin[fi]; ap < part.app_end[fi]; ap++) {
          for (int e = 0; e < E; e++) {
            const int d = prob.kkt_dest[(size_t) (prob.trip_base[fi] + (long long) ap * E + e)];
            if (d >= 0) touch(OA_KKT, d);
          }
          if (fr.role != 0)
            for (int o = 0; o < fr.orows; o++) {
              const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) ap * fr.orows + o)];
              if (d >= 0) touch(fr.role == 1 ? OA_FXE : OA_FXI, d);
            }
        }
      }
      for (int a = 0; a < OA_COUNT; a++)
        for (int w = 0; w < 2; w++)
          for (int t: part.g[a][w].target) touch(a, t);
    }
    bool ok = true;
    for (int a = 0; a < OA_COUNT && ok; a++) {
      std::vector<unsigned char> flag(part.cut_target[a].size());
      for (size_t q = 0; q < flag.size(); q++)
        flag[q] = (unsigned char) ((part.cut_mine[a][q] ? 1 : 0) | (part.cut_owner[a][q] == rank ? 2 : 0));
      d_cut_target[a] = dupload(part.cut_target[a]);
      d_cut_flag[a] = dupload(flag);
      ok = d_cut_target[a] && d_cut_flag[a];
    }
    if (!ok) {
      last_error = "out of device memory for the cut set";
      return (int) cudaErrorMemoryAllocation;
    }
  }
  // Staging slots that only ever receive a structural zero (the specialised kernels do not store those, jit.cpp) are
  // removed from the gather lists: the staging buffer is zero-filled above and the sums lose nothing.
  GatherList pruned[OA_COUNT][2], pruned_full[OA_COUNT][2];
  if (use_jit && !getenv("ASSET_B200_NO_PRUNE")) {
    std::vector<unsigned char> dead((size_t) std::max<long long>(prob.stage_size, 1), 0);
    for (size_t k = 0; k < fns.size(); k++) {
      const int fi = prob.order[k];
      const FuncRec& fr = prob.funcs[fi];
      const int E = fr.triplets_per_app();
      for (const auto& z: const_zero_outputs(host_progs[k][FM_ALL])) {
        for (int ap = 0; ap < fr.nappl; ap++) {
          long long slot = -1;
          if (z.first == PO_H || z.first == PO_J) {
            const int d = prob.kkt_dest[(size_t) (prob.trip_base[fi] + (long long) ap * E + z.second)];
            if (d < 0) slot = ~d;
          } else if (z.first == PO_G) {
            slot = prob.g_stage[fi] + (long long) ap * fr.ir + z.second;
          } else if (z.first == PO_F && fr.role != ROLE_OBJ) {
            const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) ap * fr.orows + z.second)];
            if (d < 0) slot = ~d;
          } else if (z.first == PO_F) {
            slot = prob.v_stage[fi] + ap;
          }
          if (slot >= 0) dead[(size_t) slot] = 1;
        }
      }
    }
    for (int a = 0; a < OA_COUNT; a++) {
      for (int w = 0; w < 2; w++) pruned[a][w] = prune_gather(gsrc[a][w], dead);
      gsrc[a] = pruned[a];
    }
    if (world > 1) {  // the fused evaluation of a sharded plan starts from the unpartitioned lists
      const GatherList* full[OA_COUNT] = {prob.g_kkt, prob.g_pgx, prob.g_agx, prob.g_fxe, prob.g_fxi, prob.g_val};
      for (int a = 0; a < OA_COUNT; a++)
        for (int w = 0; w < 2; w++) pruned_full[a][w] = prune_gather(full[a][w], dead);
    }
  }
  if (const char* e = getenv("ASSET_B200_FUSED")) use_fused = std::atoi(e) != 0;
  if (use_fused && use_jit && (world == 1 || !getenv("ASSET_B200_NO_PRUNE"))) {
    fused = build_fused(prob, jit_min_apps);
    if (fused.ok) {
      const long long R = fused.t_hi - fused.t_lo;
      // staging slots the fused kernels write (objective value per segment, entries of rows no segment owns)
      std::vector<unsigned char> fslot;
      {
        fslot.assign((size_t) std::max<long long>(prob.stage_size, 1), 0);
        for (const auto& pc: fused.pieces)
          for (const JitOut& w: pc.fout)
            if (w.arr == FA_STAGE)  // 2: written by this rank's segments, 1: by another rank's
              for (long long r = 0; r < R; r++)
                fslot[(size_t) (w.base + r * w.stride)] = (r >= R * rank / world && r < R * (rank + 1) / world) ? 2 : 1;
      }
      // Gather lists of the fused evaluation: targets inside the fused rows are gone.  Sharded plans: rank 0 runs the
      // per-function path of the mesh ends and owns every remaining target; the other ranks keep only the targets fed
      // by their fused kernels' staged outputs (without the solver coefficients, which rank 0 adds) and all ranks add
      // those up through the cut buffer.
      const GatherList* base[OA_COUNT];
      for (int a = 0; a < OA_COUNT; a++) base[a] = world > 1 ? pruned_full[a] : gsrc[a];
      const long long own_lo[OA_COUNT] = {fused.kb, fused.vb, fused.vb, 0, 0, 0};
      const long long own_hi[OA_COUNT] = {fused.kb + R * fused.ks, fused.vb + R * fused.vs, fused.vb + R * fused.vs, 0, 0, 0};
      // fl: targets that must wait for the fused kernels (fed by their staged outputs) and every high fan-in target;
      // flE ("early"): low fan-in targets fed by the mesh ends and the solver coefficients alone -- most of what is
      // left, the pivot and slack rows -- which are summed on the side stream beside the bulk kernels
      // flS: KKT targets fed by solver coefficients alone (pivot, slack and diagonal entries outside the fused rows): they
      // change only when the coefficients do, so set_solver_coeffs() writes them and an evaluation leaves them alone
      GatherList fl[OA_COUNT][2], flE[OA_COUNT][2], flS[2];
      for (int w = 0; w < 2; w++) flS[w].ptr.assign(1, 0);
      std::vector<int> cut_t[OA_COUNT];
      for (int a = 0; a < OA_COUNT; a++) {
        for (int w = 0; w < 2; w++) {
          const GatherList& g = base[a][w];
          fl[a][w].ptr.assign(1, 0);
          flE[a][w].ptr.assign(1, 0);
          // solver-only targets of this list (outside the fused rows), for an even spread over the ranks
          long long total_solver_only = 0, n_solver_only = 0;
          if (world > 1)
            for (int t = 0; t < g.ntargets(); t++) {
              if (g.target[t] >= own_lo[a] && g.target[t] < own_hi[a]) continue;
              bool so = g.ptr[(size_t) t + 1] > g.ptr[t];
              for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1] && so; k++) {
                const int sidx = g.src[(size_t) k];
                if (sidx < prob.stage_solver || sidx >= prob.stage_solver + prob.n_solver) so = false;
              }
              if (so) total_solver_only++;
            }
          for (int t = 0; t < g.ntargets(); t++) {
            if (g.target[t] >= own_lo[a] && g.target[t] < own_hi[a]) continue;
            bool fed = false;
            for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1]; k++) fed = fed || fslot[(size_t) g.src[(size_t) k]] != 0;
            const bool early = !fed && w == 0 && (a == OA_KKT || a == OA_PGX || a == OA_AGX);
            bool sonly = early && a == OA_KKT && world == 1 && g.ptr[(size_t) t + 1] > g.ptr[t];
            for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1] && sonly; k++) {
              const int sidx = g.src[(size_t) k];
              if (sidx < prob.stage_solver || sidx >= prob.stage_solver + prob.n_solver) sonly = false;
            }
            GatherList& o = sonly ? flS[w] : early ? flE[a][w] : fl[a][w];
            // targets fed only by solver coefficients (pivot and slack rows: most of what is left) are spread over the
            // ranks by position; everything else that no fused kernel feeds belongs to rank 0, which runs the mesh ends
            bool solver_only = world > 1 && !fed && g.ptr[(size_t) t + 1] > g.ptr[t];
            if (solver_only)
              for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1]; k++) {
                const int sidx = g.src[(size_t) k];
                if (sidx < prob.stage_solver || sidx >= prob.stage_solver + prob.n_solver) solver_only = false;
              }
            const int towner = solver_only ? (int) (n_solver_only++ * world / std::max<long long>(total_solver_only, 1)) : 0;
            if (world > 1 && !fed && rank != towner) continue;
            for (long long k = g.ptr[t]; k < g.ptr[(size_t) t + 1]; k++) {
              const int sidx = g.src[(size_t) k];
              if (world > 1 && !solver_only) {
                // a rank sums what it wrote in this evaluation: its own segments' staged outputs, and on rank 0 the mesh
                // ends' and the solver coefficients (slots of other modes' evaluations may hold stale values)
                const unsigned char fs = fslot[(size_t) sidx];
                if (fs == 1 || (rank != 0 && fs != 2)) continue;
              }
              o.src.push_back(sidx);
            }
            o.ptr.push_back((long long) o.src.size());
            o.target.push_back(g.target[t]);
            if (fed) cut_t[a].push_back(g.target[t]);
          }
        }
      }
      if (windowed) {
        try {
          // the solver-only entries go through the same ownership rules
          for (int w = 0; w < 2; w++) {
            GatherList& e = flE[OA_KKT][w];
            const GatherList& q = flS[w];
            for (int t = 0; t < q.ntargets(); t++) {
              e.src.insert(e.src.end(), q.src.begin() + q.ptr[(size_t) t], q.src.begin() + q.ptr[(size_t) t + 1]);
              e.ptr.push_back((long long) e.src.size());
              e.target.push_back(q.target[(size_t) t]);
            }
            flS[w] = GatherList();
            flS[w].ptr.assign(1, 0);
          }
          window_filter(fl, flE, cut_t, true_world);
          // ... and come out again: owned entries fed by solver coefficients alone
          for (int w = 0; w < 2; w++) {
            GatherList keep;
            keep.ptr.assign(1, 0);
            GatherList& e = flE[OA_KKT][w];
            const auto& cut = cut_t[OA_KKT];
            for (int t = 0; t < e.ntargets(); t++) {
              bool so = e.ptr[(size_t) t + 1] > e.ptr[(size_t) t] && !std::binary_search(cut.begin(), cut.end(), e.target[(size_t) t]);
              for (long long k = e.ptr[(size_t) t]; k < e.ptr[(size_t) t + 1] && so; k++)
                if (e.src[(size_t) k] < prob.stage_solver || e.src[(size_t) k] >= prob.stage_solver + prob.n_solver) so = false;
              GatherList& o = so ? flS[w] : keep;
              o.src.insert(o.src.end(), e.src.begin() + e.ptr[(size_t) t], e.src.begin() + e.ptr[(size_t) t + 1]);
              o.ptr.push_back((long long) o.src.size());
              o.target.push_back(e.target[(size_t) t]);
            }
            e = std::move(keep);
          }
        } catch (const std::exception& ex) {
          last_error = ex.what();
          return (int) cudaErrorInvalidValue;
        }
      }
      std::vector<double> appc;
      for (const FuncRec& fr: prob.funcs) appc.insert(appc.end(), fr.appconst.begin(), fr.appconst.end());
      d_appc_all = dupload(appc);
      bool ok = d_appc_all && up_gather(*this, fl[OA_KKT], gkF) && up_gather(*this, fl[OA_PGX], gpF) && up_gather(*this, fl[OA_AGX], gaF)
                && up_gather(*this, flE[OA_KKT], gkE) && up_gather(*this, flE[OA_PGX], gpE) && up_gather(*this, flE[OA_AGX], gaE)
                && up_gather(*this, flS, gkS);
      if (ok && windowed) {
        ok = up_gather(*this, fl[OA_FXE], geF) && up_gather(*this, fl[OA_FXI], giF) && up_gather(*this, fl[OA_VAL], gvF);
        cutF_count = 0;
        for (int a = 0; a < OA_COUNT && ok; a++) {
          cutF_n[a] = (long long) cut_t[a].size();
          cutF_count += cutF_n[a];
          std::vector<unsigned char> flag(cut_t[a].size(), (unsigned char) (rank == 0 ? 3 : 1));
          d_cutF_target[a] = dupload(cut_t[a]);
          d_cutF_flag[a] = dupload(flag);
          ok = d_cutF_target[a] && d_cutF_flag[a];
        }
      }
      if (ok && world > 1) {
        ok = up_gather(*this, fl[OA_FXE], geF) && up_gather(*this, fl[OA_FXI], giF) && up_gather(*this, fl[OA_VAL], gvF);
        cutF_count = 0;
        for (int a = 0; a < OA_COUNT && ok; a++) {
          cutF_n[a] = (long long) cut_t[a].size();
          cutF_count += cutF_n[a];
          std::vector<unsigned char> flag(cut_t[a].size(), (unsigned char) (rank == 0 ? 3 : 1));
          d_cutF_target[a] = dupload(cut_t[a]);
          d_cutF_flag[a] = dupload(flag);
          ok = d_cutF_target[a] && d_cutF_flag[a];
        }
        // what this rank writes in the fused evaluation: its segments' rows, their constraint rows and far entries;
        // rank 0 also everything the mesh ends and the remaining gather targets touch
        const long long r0 = R * rank / world, r1 = R * (rank + 1) / world;
        auto touch = [&](int a, long long i) { out_blocks[a][(size_t) (i / OUT_BLOCK)] = 1; };
        auto arr_of = [](int fa) {
          return fa == FA_KKT ? (int) OA_KKT : fa == FA_FXE ? (int) OA_FXE : fa == FA_FXI ? (int) OA_FXI
                 : fa == FA_PGX ? (int) OA_PGX : fa == FA_AGX ? (int) OA_AGX : -1;
        };
        for (const auto& pc: fused.pieces)
          for (const JitOut& w: pc.fout) {
            const int a = arr_of(w.arr);
            if (a < 0 || r1 <= r0) continue;
            const long long i0 = w.base + r0 * w.stride, i1 = w.base + (r1 - 1) * w.stride;
            for (long long i = std::min(i0, i1) / OUT_BLOCK; i <= std::max(i0, i1) / OUT_BLOCK; i++) out_blocks[a][(size_t) i] = 1;
          }
        for (int a = 0; a < OA_COUNT; a++)
          for (int w = 0; w < 2; w++)
          {
            for (int t: fl[a][w].target) touch(a, t);
            for (int t: flE[a][w].target) touch(a, t);
          }
        if (rank == 0) {
          std::vector<int> clr[OA_COUNT];
          for (const auto& pc: fused.pieces)
            for (const JitOut& w: pc.fout) {
              const int a = arr_of(w.arr);
              if (a < 0) continue;
              for (long long r = std::max(r1, R - 4); r < R; r++) clr[a].push_back((int) (w.base + r * w.stride));
            }
          for (int a = 0; a < OA_COUNT && ok; a++) {
            clear_n[a] = (long long) clr[a].size();
            d_clear[a] = dupload(clr[a]);
            ok = d_clear[a] != nullptr;
          }
          for (int fi: prob.order) {
            const FuncRec& fr = prob.funcs[fi];
            const int E = fr.triplets_per_app();
            for (int ap = 0; ap < fr.nappl; ap++) {
              if (ap >= fused.ex_lo[fi] && ap < fused.ex_hi[fi]) {
                ap = fused.ex_hi[fi] - 1;
                continue;
              }
              for (int e = 0; e < E; e++) {
                const int d = prob.kkt_dest[(size_t) (prob.trip_base[fi] + (long long) ap * E + e)];
                if (d >= 0) touch(OA_KKT, d);
              }
              if (fr.role != 0)
                for (int o = 0; o < fr.orows; o++) {
                  const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) ap * fr.orows + o)];
                  if (d >= 0) touch(fr.role == 1 ? OA_FXE : OA_FXI, d);
                }
            }
          }
        }
      }
      // destination maps of the applications left to the per-function kernels, fused-owned slots redirected to the trash
      // slot (see plan.hpp: kdF)
      if (ok) {
        const long long own_k0 = fused.kb, own_k1 = fused.kb + R * fused.ks;
        const int trash = ~(int) prob.stage_size;
        kdF.assign(fns.size(), nullptr);
        kdF_lo.assign(fns.size(), 0);
        kdF_hi.assign(fns.size(), 0);
        for (size_t k = 0; k < fns.size() && ok; k++) {
          const int fi = prob.order[k];
          const FuncRec& fr = prob.funcs[fi];
          const int E = fr.triplets_per_app();
          int xl = fused.ex_lo[fi], xh = fused.ex_hi[fi];
          if (xh <= xl) xl = xh = fr.nappl;  // not covered by the fused kernels: every application, one chunk
          const long long rows = (long long) xl + (fr.nappl - xh);
          if (rows > (1 << 20)) {  // a large function outside the template: keep its own map (it cannot touch fused rows)
            kdF_lo[k] = kdF_hi[k] = -1;
            continue;
          }
          std::vector<int> m((size_t) (rows * E));
          const int* src = prob.kkt_dest.data() + prob.trip_base[fi];
          for (long long r = 0; r < rows; r++) {
            const long long a = r < xl ? r : xh + (r - xl);
            for (int e = 0; e < E; e++) {
              const int d = src[(size_t) (a * E + e)];
              m[(size_t) (r * E + e)] = (d >= own_k0 && d < own_k1) ? trash : d;
            }
          }
          if (m.empty()) m.push_back(0);
          kdF[k] = dupload(m);
          kdF_lo[k] = xl;
          kdF_hi[k] = xh;
          ok = kdF[k] != nullptr;
        }
      }
      if (!ok) {
        last_error = "out of device memory for the fused evaluation's gather lists";
        return (int) cudaErrorMemoryAllocation;
      }
    } else {
      jit_log = "no fused interior: " + fused.why;
    }
  }
  if (windowed) {
    if (!fused.ok) {
      last_error = "a windowed plan needs the fused evaluation: " + (fused.why.empty() ? jit_log : fused.why);
      return (int) cudaErrorInvalidValue;
    }
    world = true_world;
    d_cutbuf = dalloc<double>((size_t) std::max<long long>(cutF_count, 1));
    if (!d_cutbuf) {
      last_error = "out of device memory for the cut buffer";
      return (int) cudaErrorMemoryAllocation;
    }
    finalize_out_ranges();
    release_windowed = true;
  } else if (world > 1) {
    d_cutbuf = dalloc<double>((size_t) std::max<long long>(std::max(cut_count, cutF_count), 1));
    if (!d_cutbuf) {
      last_error = "out of device memory for the cut buffer";
      return (int) cudaErrorMemoryAllocation;
    }
    finalize_out_ranges();
  }
  if (!up_gather(*this, gsrc[OA_KKT], gk) || !up_gather(*this, gsrc[OA_PGX], gp) || !up_gather(*this, gsrc[OA_AGX], ga)
      || !up_gather(*this, gsrc[OA_FXE], ge) || !up_gather(*this, gsrc[OA_FXI], gi) || !up_gather(*this, gsrc[OA_VAL], gv)) {
    last_error = "out of device memory for the gather lists";
    return (int) cudaErrorMemoryAllocation;
  }
  {
    size_t need = 1;
    for (const DevGather* g: {&gk, &gp, &ga, &ge, &gi, &gv, &gkF, &gpF, &gaF, &geF, &giF, &gvF, &gkE, &gpE, &gaE, &gkS})
      need = std::max(need, (size_t) g->dev[1].ntargets * (size_t) g->dev[1].maxparts);
    d_gscratch = dalloc<double>(need);
    if (!d_gscratch) {
      last_error = "out of device memory for the gather scratch";
      return (int) cudaErrorMemoryAllocation;
    }
  }
  for (int w = 0; w < 3; w++) in_ranges[w] = prob.input_ranges(w);
  if (release_windowed && !getenv("ASSET_B200_KEEP_HOST_ANALYSIS")) {
    // assemble the fused kernels now (a piece that needs more shared memory than a block has re-cuts the rows, which
    // reads the analysis), then let the analysis go
    const int fr = ensure_fused(0);
    if (fr < 0) return -fr;
    if (fr == 0) {
      last_error = "a windowed plan needs the fused kernels: " + jit_log;
      return (int) cudaErrorInvalidValue;
    }
    release_host_analysis();
  }
  AB_CUDA(cudaDeviceSynchronize());
  return 0;
}

int Plan::set_solver_coeffs(const double* host) {
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(cudaMemcpyAsync(d_stage + prob.stage_solver, host, sizeof(double) * (size_t) prob.n_solver,
                          cudaMemcpyHostToDevice, stream));
  solver_entries_valid = false;  // the next fused evaluation rewrites the entries that hold these coefficients alone
  AB_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

int Plan::prof_begin(const std::string& label) {
  if (!profiling) return 0;
  const size_t k = prof_labels.size();
  while (prof_events.size() < 2 * (k + 1)) {
    cudaEvent_t e;
    AB_CUDA(cudaEventCreate(&e));
    prof_events.push_back(e);
  }
  prof_labels.push_back(label);
  AB_CUDA(cudaEventRecord(prof_events[2 * k], stream));
  return 0;
}
int Plan::prof_end() {
  if (!profiling) return 0;
  AB_CUDA(cudaEventRecord(prof_events[2 * (prof_labels.size() - 1) + 1], stream));
  return 0;
}
int Plan::profile_read(float* ms, int cap) {
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(cudaStreamSynchronize(stream));
  const int n = (int) prof_labels.size();
  for (int k = 0; k < n && k < cap; k++) AB_CUDA(cudaEventElapsedTime(&ms[k], prof_events[2 * k], prof_events[2 * k + 1]));
  return n;
}

// Process-wide cache of assembled kernels keyed by their PTX text: plans of the same problem (tests, mesh iterations
// that keep N) share the cubin instead of assembling it again.
namespace {
struct CubinEntry {
  std::vector<char> cubin;
  int nregs = 0, spill = 0;
  unsigned smem = 0;  // fused kernels: dynamic shared memory of the variant that was kept (0: as emitted)
};
std::mutex g_cubin_mu;
std::unordered_map<std::string, CubinEntry> g_cubins;
}  // namespace

int Plan::ensure_fused(int mode) {
  // mode 0: the evalKKT pieces; 4..7: the residual-only program of that evaluation mode (fused.hpp)
  signed char& fused_state = mode == 0 ? this->fused_state : fused_res_state[mode & 7];
  std::vector<JitKernel>& fused_jits = mode == 0 ? this->fused_jits : fused_res_jits[mode & 7];
  const std::vector<FusedPlan::Piece>& pieces = mode == 0 ? fused.pieces : fused.residual[mode & 7];
  if (fused_state != 0) return fused_state > 0 ? 1 : 0;
  fused_state = -1;
  if (!fused.ok || pieces.empty()) return 0;
  int dev_smem = 0;
  cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  fused_jits.assign(pieces.size(), JitKernel());
  for (size_t pc = 0; pc < pieces.size(); pc++) {
    std::string ptx;
    std::vector<JitOut> outs;
    JitStats js;
    try {
      ptx = fused_ptx(fused, (int) pc, &outs, &js, 0, mode);
    } catch (const std::exception& e) {
      jit_log = std::string("fused kernel: ") + e.what();
      return 0;
    }
    if ((int) js.smem_bytes > dev_smem) {
      // too many values alive for one block: cut the rows into more pieces and start over
      const int np = (int) fused.pieces.size() + 1;
      if (mode != 0 || np > 8 || np > fused.vs) {
        jit_log = "fused kernel needs " + std::to_string(js.smem_bytes) + " B of shared memory per block";
        return 0;  // the per-function path evaluates everything
      }
      for (JitKernel& k: fused_jits) unload_jit(k);
      fused = build_fused(prob, jit_min_apps, np);
      if (!fused.ok) return 0;
      fused_state = 0;
      return ensure_fused(0);
    }
    CubinEntry ce;
    bool have = false;
    {
      std::lock_guard<std::mutex> lk(g_cubin_mu);
      auto it = g_cubins.find(ptx);
      if (it != g_cubins.end()) {
        ce = it->second;
        have = true;
      }
    }
    if (!have) {
      std::string log;
      if (!assemble_ptx(ptx, ce.cubin, log, &ce.nregs, &ce.spill)) {
        last_error = "fused kernel: " + log;
        return -(int) cudaErrorInvalidPtx;
      }
      // The assembler's own spills go to local memory, whose reloads queue behind the output stream (ncu, capture s2: a
      // reload of a spilled address pair waited ~2500 cycles): give it room by keeping fewer values in registers.
      const std::string first = ptx;
      int budget = js.reg_budget;
      ce.smem = js.smem_bytes;
      for (int tries = 0; ce.spill > 0 && tries < 5 && !getenv("ASSET_B200_FUSED_KEEP_SPILLS"); tries++) {
        JitStats s2;
        budget -= 3;
        if (budget < 48) break;
        const std::string p2 = fused_ptx(fused, (int) pc, nullptr, &s2, budget, mode);
        if ((int) s2.smem_bytes > dev_smem) break;
        CubinEntry c2;
        if (!assemble_ptx(p2, c2.cubin, log, &c2.nregs, &c2.spill)) break;
        if (c2.spill < ce.spill) {
          ce = c2;
          ce.smem = s2.smem_bytes;
        }
      }
      std::lock_guard<std::mutex> lk(g_cubin_mu);
      g_cubins.emplace(first, ce);
    }
    JitKernel& jk = fused_jits[pc];
    jk.smem_bytes = ce.smem ? ce.smem : js.smem_bytes;
    cudaError_t e = load_jit(ce.cubin, mode == 0 ? "ab_fused_kkt" + std::to_string(pc) : "ab_fused_m" + std::to_string(mode) + "_" + std::to_string(pc), jk);
    if (e != cudaSuccess) {
      last_error = std::string("loading the fused kernel: ") + cudaGetErrorString(e);
      return -(int) e;
    }
    const size_t n = (outs.size() + 31) / 32 * 32;
    std::vector<long long> tb(n, 0);
    std::vector<int> ts(n, 0);
    for (size_t q = 0; q < outs.size(); q++) {
      const JitOut& w = outs[q];
      double* arr = nullptr;
      switch (w.arr) {
        case FA_KKT: arr = d_kkt; break;
        case FA_FXE: arr = d_fxe; break;
        case FA_FXI: arr = d_fxi; break;
        case FA_STAGE: arr = d_stage; break;
        case FA_PGX: arr = d_pgx; break;
        case FA_AGX: arr = d_agx; break;
        default: last_error = "fused kernel: unknown output array"; return -(int) cudaErrorInvalidValue;
      }
      tb[q] = (long long) (uintptr_t) arr + w.base * 8;
      ts[q] = (int) (w.stride * 8);
    }
    jk.out_base = dupload(tb);
    jk.out_stride = dupload(ts);
    if (!jk.out_base || !jk.out_stride) {
      last_error = "out of device memory for the output table of the fused kernel";
      return -(int) cudaErrorMemoryAllocation;
    }
    jk.nthreads = 32;
    if (const char* env = getenv("ASSET_B200_JIT_THREADS")) jk.nthreads = std::max(32, std::atoi(env) / 32 * 32);
    if (const char* env = getenv("ASSET_B200_FUSED_THREADS")) jk.nthreads = std::max(32, std::atoi(env) / 32 * 32);
    jk.nregs = ce.nregs;
    jk.spill_bytes = ce.spill;
    if (getenv("ASSET_B200_JIT_VERBOSE"))
      fprintf(stderr, "[asset_b200] ab_fused_kkt%zu of %zu: segments [%d, %d), %zu application families, %zu loads, %zu stored "
                      "outputs, %d FMA, %d regs, %d B assembler spill, %u B smem/block, %d parked slots, cubin %zu B\n",
              pc, pieces.size(), fused.t_lo, fused.t_hi, fused.insts.size(), fused.fin.size(), outs.size(),
              pieces[pc].prog.n_fma, ce.nregs, ce.spill, jk.smem_bytes, js.parked_slots, ce.cubin.size());
  }
  fused_state = 1;
  return 1;
}

int Plan::ensure_jit(size_t k, int fm, int variant) {
  if (!use_jit) return 0;
  signed char& st = jit_state[variant][k][fm];
  if (st != 0) return st > 0 ? 1 : 0;
  st = -1;
  const AffineMaps& am = affs[k];
  if (variant == 0 && am.a1 - am.a0 < jit_min_apps) return 0;
  const FuncRec& fr = prob.funcs[prob.order[k]];
  JitSpec sp;
  sp.name = std::string(variant ? "ab_idx_f" : "ab_seg_f") + std::to_string(k) + "_m" + std::to_string(fm);
  sp.role = fr.role;
  sp.ir = fr.ir;
  sp.orows = fr.orows;
  sp.nappconst = fr.nappconst;
  sp.E = fr.triplets_per_app();
  sp.indexed = variant == 1;
  sp.g_stage = fns[k].g_stage;
  sp.v_stage = fns[k].v_stage;
  sp.nthreads = variant ? 32 : 64;
  std::string ptx;
  std::vector<JitOut> outs;
  JitStats js;
  jit_env_knobs(sp);
  if (variant == 1 && getenv("ASSET_B200_IDX_NO_PARK")) sp.park_indexed = false;
  try {
    ptx = emit_ptx(host_progs[k][fm], sp, am, &outs, &js);
    int dev_smem = 0;
    cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (variant == 1 && (int) js.smem_bytes > dev_smem) {  // more live values than one block can park: leave them to the assembler
      sp.park_indexed = false;
      ptx = emit_ptx(host_progs[k][fm], sp, am, &outs, &js);
    }
  } catch (const std::exception& e) {
    jit_log = std::string(e.what()) + " (" + fr.name + ")";
    return 0;
  }
  CubinEntry ce;
  bool have = false;
  {
    std::lock_guard<std::mutex> lk(g_cubin_mu);
    auto it = g_cubins.find(ptx);
    if (it != g_cubins.end()) {
      ce = it->second;
      have = true;
    }
  }
  if (!have) {
    std::string log;
    if (!assemble_ptx(ptx, ce.cubin, log, &ce.nregs, &ce.spill)) {
      // a program the assembler rejects is a bug of the emitter: fail loudly rather than fall back silently
      last_error = "specialised kernel of " + fr.name + ": " + log;
      return -(int) cudaErrorInvalidPtx;
    }
    std::lock_guard<std::mutex> lk(g_cubin_mu);
    g_cubins.emplace(ptx, ce);
  }
  JitKernel& jk = jits[variant][k][fm];
  jk.smem_bytes = js.smem_bytes;
  cudaError_t e = load_jit(ce.cubin, sp.name, jk);
  if (e != cudaSuccess) {
    last_error = "loading the specialised kernel of " + fr.name + ": " + cudaGetErrorString(e);
    return -(int) e;
  }
  if (variant == 0) {
    // absolute destination table of the cooperative stores (padded to whole 32-output tiles)
    const size_t n = (outs.size() + 31) / 32 * 32;
    std::vector<long long> tb(n, 0);
    std::vector<int> ts(n, 0);
    for (size_t q = 0; q < outs.size(); q++) {
      const JitOut& w = outs[q];
      double* arr = (w.arr == 2) ? d_kkt : (w.arr == 3) ? ((fr.role == ROLE_EQ) ? d_fxe : d_fxi) : d_stage;
      tb[q] = (long long) (uintptr_t) arr + w.base * 8;
      ts[q] = (int) (w.stride * 8);
    }
    jk.out_base = dupload(tb);
    jk.out_stride = dupload(ts);
    if (!jk.out_base || !jk.out_stride) {
      last_error = "out of device memory for the output table of " + fr.name;
      return -(int) cudaErrorMemoryAllocation;
    }
  }
  if (getenv("ASSET_B200_JIT_VERBOSE"))
    fprintf(stderr, "[asset_b200] %s (%s): %d regs, %d B assembler spill, %u B smem/block, %d parked slots, %d parks, %d reloads, cubin %zu B\n",
            sp.name.c_str(), fr.name.c_str(), ce.nregs, ce.spill, js.smem_bytes, js.parked_slots, js.parks, js.reloads,
            ce.cubin.size());
  jk.nthreads = sp.nthreads;
  jk.nregs = ce.nregs;
  jk.spill_bytes = ce.spill;
  st = 1;
  return 1;
}

int Plan::eval(int mode, double obj_scale) {
  AB_CUDA(cudaSetDevice(device));
  static const bool graphs_off = getenv("ASSET_B200_GRAPH") && std::atoi(getenv("ASSET_B200_GRAPH")) == 0;
  static const bool timeline_on = getenv("ASSET_B200_TIMELINE") != nullptr;
  const bool fused_now = mode == EM_KKT && fused.ok && fused_state >= 0;
  if (fused_now && !solver_entries_valid && gkS.dev[0].ntargets + gkS.dev[1].ntargets > 0) {
    // entries of the value array that hold solver coefficients alone (set_solver_coeffs): not part of an evaluation
    AB_CUDA(launch_gather(gkS.dev[0], gkS.dev[1], d_stage, d_kkt, d_gscratch, stream));
    solver_entries_valid = true;
  }
  // a sharded plan zero-fills its outputs when the evaluation path changes: that call is not the one to capture or replay
  const bool transition = world > 1 && !windowed && fused_now != last_fused;
  if (!use_graphs || graphs_off || timeline_on || profiling || mode < 0 || mode >= 8 || transition) return enqueue(mode, obj_scale);
  EvalGraph& g = graphs[mode];
  if (g.failed) return enqueue(mode, obj_scale);
  if (g.exec && g.obj_scale == obj_scale) {
    AB_CUDA(cudaGraphLaunch(g.exec, stream));
    launches = g.launches;
    jit_launches = g.jit_launches;
    last_mode = mode;
    last_fused = g.fused;
    return 0;
  }
  if (g.evals < 1) {  // first evaluation of the mode: kernels are assembled and tables uploaded on the way
    g.evals++;
    return enqueue(mode, obj_scale);
  }
  if (g.exec) {
    cudaGraphExecDestroy(g.exec);
    g.exec = nullptr;
  }
  cudaGraph_t graph = nullptr;
  if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    g.failed = true;
    return enqueue(mode, obj_scale);
  }
  const int rc = enqueue(mode, obj_scale);
  const cudaError_t ce = cudaStreamEndCapture(stream, &graph);
  if (rc != 0 || ce != cudaSuccess || !graph) {
    cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    g.failed = true;
    return rc != 0 ? rc : enqueue(mode, obj_scale);
  }
  const cudaError_t ie = cudaGraphInstantiate(&g.exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ie != cudaSuccess) {
    cudaGetLastError();
    g.exec = nullptr;
    g.failed = true;
    return enqueue(mode, obj_scale);
  }
  g.obj_scale = obj_scale;
  g.fused = last_fused;
  g.launches = launches;
  g.jit_launches = jit_launches;
  AB_CUDA(cudaGraphLaunch(g.exec, stream));
  return 0;
}

int Plan::enqueue(int mode, double obj_scale) {
  int obj_mode = -1, con_mode = -1;
  bool out_kkt = false, out_pgx = false, out_agx = false, out_fx = false, out_val = false;
  switch (mode) {
    case EM_KKT: obj_mode = FM_ALL; con_mode = FM_ALL; out_kkt = out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_KKTNO: con_mode = FM_ALL; out_kkt = out_agx = out_fx = true; break;
    case EM_SOE: con_mode = FM_JAC; out_kkt = out_fx = true; break;
    case EM_AUG: obj_mode = FM_GRAD; con_mode = FM_JAC_GRAD; out_kkt = out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_RHS: obj_mode = FM_GRAD; con_mode = FM_GRAD; out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_OGC: obj_mode = FM_GRAD; con_mode = FM_VALUES; out_pgx = out_fx = out_val = true; break;
    case EM_OCC: obj_mode = FM_VALUES; con_mode = FM_VALUES; out_fx = out_val = true; break;
    case EM_OBJ: obj_mode = FM_VALUES; out_val = true; break;
    default: last_error = "unknown eval mode"; return (int) cudaErrorInvalidValue;
  }
  if (windowed && mode != EM_KKT) {
    last_error = "a windowed plan evaluates evalKKT only";
    return (int) cudaErrorNotSupported;
  }
  launches = 0;
  jit_launches = 0;
  last_mode = mode;
  if (profiling != 2) prof_labels.clear();
  if (out_kkt && mode != EM_KKT) {
    // partial matrix modes: entries this mode does not compute must read as zero
    AB_CUDA(cudaMemsetAsync(d_kkt, 0, sizeof(double) * (size_t) prob.nnz(), stream));
    AB_CUDA(cudaMemsetAsync(d_stage, 0, sizeof(double) * (size_t) prob.stage_solver, stream));
    solver_entries_valid = false;
  }
  // Every function writes its own outputs (shared entries go through the staging buffer), so the launches are
  // independent: the bulk kernels run back to back on `stream`, the small ones (mesh ends, functions with a handful of
  // applications) overlap them on `side`.  With per-launch profiling on everything is serialised on `stream`.
  const bool overlap = !profiling;
  bool forked = false, forked2 = false, joined = false;
  // the forked streams are joined on every way out of this function: an error return between fork and join must not
  // leave side-stream work running beside the next evaluation
  unsigned pool_used = 0;  // side_pool streams (beyond `side`) that carry work of this evaluation
  int pool_next = 0;
  struct Joiner {
    Plan& P;
    bool &f1, &f2, &done;
    unsigned& pool;
    ~Joiner() {
      if (done) return;
      for (int i = 1; i < SIDE_POOL; i++)
        if ((pool >> i & 1u) && cudaEventRecord(P.ev_pool[i], P.side_pool[i]) == cudaSuccess) cudaStreamWaitEvent(P.stream, P.ev_pool[i], 0);
      if (f1 && cudaEventRecord(P.ev_join, P.side) == cudaSuccess) cudaStreamWaitEvent(P.stream, P.ev_join, 0);
      if (f2 && cudaEventRecord(P.ev_join2, P.bulk2) == cudaSuccess) cudaStreamWaitEvent(P.stream, P.ev_join2, 0);
    }
  } joiner{*this, forked, forked2, joined, pool_used};
  // ASSET_B200_TIMELINE=1 (measurement): device time stamps on every stream of this evaluation, printed after a
  // synchronisation -- what ran beside what
  static const bool timeline = getenv("ASSET_B200_TIMELINE") != nullptr;
  std::vector<std::pair<std::string, cudaEvent_t>> marks;
  auto mark = [&](const char* label, cudaStream_t st) {
    if (!timeline) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    marks.emplace_back(label, e);
  };
  mark("start", stream);
  auto small_stream = [&]() -> cudaStream_t {
    if (!overlap) return stream;
    if (!forked) {
      cudaEventRecord(ev_fork, stream);
      cudaStreamWaitEvent(side, ev_fork, 0);
      forked = true;
      mark("side: released", side);
    }
    return side;
  };
  // fused evaluation: the launches of the mesh ends are independent of each other and of the fused kernels (plan.hpp:
  // kdF), so they are dealt round-robin over a pool of side streams instead of queueing up on one
  auto pool_stream = [&]() -> cudaStream_t {
    if (!overlap) return stream;
    small_stream();  // forks `side` (= side_pool[0]) off the main stream on first use
    const int i = pool_next++ % SIDE_POOL;
    if (i != 0 && !(pool_used >> i & 1u)) {
      cudaStreamWaitEvent(side_pool[i], ev_fork, 0);
      pool_used |= 1u << i;
    }
    return side_pool[i];
  };
  // everything queued on the pool so far has finished when `side` gets here
  auto pool_gather = [&]() {
    if (timeline) {
      static const char* const names[SIDE_POOL] = {"pool 0: its mesh ends done", "pool 1: its mesh ends done", "pool 2: its mesh ends done",
                                                   "pool 3: its mesh ends done"};
      for (int i = 0; i < SIDE_POOL; i++)
        if (i == 0 ? forked : (pool_used >> i & 1u) != 0) mark(names[i], side_pool[i]);
    }
    for (int i = 1; i < SIDE_POOL; i++)
      if (pool_used >> i & 1u) {
        cudaEventRecord(ev_pool[i], side_pool[i]);
        cudaStreamWaitEvent(side, ev_pool[i], 0);
      }
  };
  auto bulk_stream = [&](size_t k) -> cudaStream_t {
    if (!overlap || heavy_fn < 0 || (int) k == heavy_fn) return stream;
    if (!forked2) {
      if (!forked) cudaEventRecord(ev_fork, stream);
      cudaStreamWaitEvent(bulk2, ev_fork, 0);
      forked2 = true;
    }
    return bulk2;
  };
  bool fused_eval = false;
  // the fused kernels evaluate evalKKT and -- single-device plans -- the residual-only modes of the line search
  const bool residual_mode = mode >= EM_RHS && mode <= EM_OBJ && world == 1 && !windowed && !getenv("ASSET_B200_NO_FUSED_RESIDUAL");
  // The partial-matrix evaluations (evalKKTNO / SOE / AUG) on the fused threads: opt-in (ASSET_B200_FUSED_PARTIAL=1) until
  // they have been run on a device -- written after the round's GPU time was spent; the kernels are checked on the CPU
  // (tests/test_ptx_emulation.py), the GPU tests are tests/test_zz3_gpu_partial_modes.py.
  const bool partial_mode = mode >= EM_KKTNO && mode <= EM_AUG && world == 1 && !windowed && getenv("ASSET_B200_FUSED_PARTIAL") != nullptr
                            && std::atoi(getenv("ASSET_B200_FUSED_PARTIAL")) != 0;
  if ((mode == EM_KKT || residual_mode || partial_mode) && fused.ok) {
    const int fr = ensure_fused(mode == EM_KKT ? 0 : mode);
    if (fr < 0) return -fr;
    fused_eval = fr > 0;
  }
  if (fused_eval && partial_mode && gkS.dev[0].ntargets + gkS.dev[1].ntargets > 0) {
    // the entries that hold solver coefficients alone are not in the fused evaluation's gather lists (an evalKKT leaves
    // them alone); the zero fill above cleared them
    AB_CUDA(launch_gather(gkS.dev[0], gkS.dev[1], d_stage, d_kkt, d_gscratch, stream));
    launches += (gkS.dev[0].ntargets > 0) + 2 * (gkS.dev[1].ntargets > 0);
    solver_entries_valid = true;
  }
  const std::vector<JitKernel>& fjits = mode == EM_KKT ? fused_jits : fused_res_jits[mode & 7];
  if (windowed && !fused_eval) {
    last_error = "a windowed plan needs the fused kernels: " + jit_log;
    return (int) cudaErrorNotSupported;
  }
  if (world > 1 && !windowed && fused_eval != last_fused) {
    // The fused evaluation and the per-function evaluation split the outputs between the ranks along slightly different
    // lines; entries this rank wrote under the other split must read as zero again ("every output is non-zero on
    // exactly one rank").
    AB_CUDA(cudaMemsetAsync(d_kkt, 0, sizeof(double) * (size_t) prob.nnz(), stream));
    solver_entries_valid = false;
    AB_CUDA(cudaMemsetAsync(d_fxe, 0, sizeof(double) * (size_t) std::max(prob.neq, 1), stream));
    AB_CUDA(cudaMemsetAsync(d_fxi, 0, sizeof(double) * (size_t) std::max(prob.niq, 1), stream));
    AB_CUDA(cudaMemsetAsync(d_pgx, 0, sizeof(double) * (size_t) std::max(prob.nvars, 1), stream));
    AB_CUDA(cudaMemsetAsync(d_agx, 0, sizeof(double) * (size_t) std::max(prob.nvars, 1), stream));
    AB_CUDA(cudaMemsetAsync(d_val, 0, sizeof(double), stream));
  }
  last_fused = fused_eval;
  for (const FnDev& f: fns) {
    int fm = (f.role == ROLE_OBJ) ? obj_mode : con_mode;
    if (fm < 0) continue;
    const size_t k = (size_t) (&f - fns.data());
    // applications this plan evaluates through the per-function kernels: its share of the function, except in the fused
    // evaluation of a sharded plan, where rank 0 alone runs what the fused kernels leave (the mesh ends)
    int fb = f.app_begin, fe = f.app_end;
    if (windowed) {  // the applications this rank owns (the fused kernels take most of them below)
      fb = prob.win.own_lo[(size_t) prob.order[k]];
      fe = prob.win.own_hi[(size_t) prob.order[k]];
    } else if (fused_eval && world > 1) {
      if (rank != 0) continue;
      fb = 0;
      fe = prob.funcs[prob.order[k]].nappl;
    }
    const std::string& name = prob.funcs[prob.order[k]].name;
    EvalIO io;
    io.X = d_X;
    io.L = (f.role == ROLE_EQ) ? d_LE : d_LI;
    io.obj_scale = obj_scale;
    io.kkt = d_kkt;
    io.fx = (f.role == ROLE_EQ) ? d_fxe : d_fxi;
    io.stage = d_stage;
    FnDev fd = f;
    fd.prog = progs[k][fm];
    if (fd.prog.nthreads == 0) {
      last_error = "no device program for this mode";
      return (int) cudaErrorInvalidValue;
    }
    JitArgs ja;
    ja.X = io.X;
    ja.L = io.L;
    ja.kkt = io.kkt;
    ja.fx = io.fx;
    ja.stage = io.stage;
    ja.appconst = f.appconst;
    ja.vindex = f.vindex;
    ja.cindex = f.cindex;
    ja.kkt_dest = f.kkt_dest;
    ja.f_dest = f.f_dest;
    ja.out_base = nullptr;
    ja.out_stride = nullptr;
    ja.obj_scale = obj_scale;
    // applications [b, e) that are not covered by the affine kernel: indexed kernel, else the interpreter
    auto rest = [&](int b, int e, const char* suffix) -> int {
      if (e <= b) return 0;
      if (int rc = prof_begin(name + suffix)) return rc;
      const int jr = ensure_jit(k, fm, 1);
      if (jr < 0) return -jr;
      cudaStream_t st = fused_eval ? pool_stream() : (e - b <= 4096) ? small_stream() : stream;
      // fused evaluation: destinations inside the fused rows are redirected to the trash slot (plan.hpp: kdF); [b, e) lies
      // in one of the two chunks of the redirected map
      const int* kd = f.kkt_dest;
      if (fused_eval && k < kdF.size() && kdF[k] && kdF_lo[k] >= 0)
        kd = (b < kdF_lo[k]) ? kdF[k] : kdF[k] + ((long long) kdF_lo[k] - kdF_hi[k]) * f.E;
      if (jr > 0) {
        JitArgs a = ja;
        a.begin = b;
        a.end = e;
        a.kkt_dest = kd;
        AB_CUDA(launch_jit(jits[1][k][fm], a, st));
        jit_launches++;
      } else {
        FnDev h = fd;
        h.app_begin = b;
        h.app_end = e;
        h.kkt_dest = kd;
        AB_CUDA(launch_eval(h, st, io));
      }
      launches++;
      return prof_end();
    };
    if (fused_eval) {
      const int fi = prob.order[k];
      const int xl = fused.ex_lo[fi], xh = fused.ex_hi[fi];
      if (xh > xl) {  // the fused kernel evaluates [xl, xh): only the applications at the ends are left
        if (int rc = rest(fb, std::min(xl, fe), ":ends")) return rc;
        if (int rc = rest(std::max(xh, fb), fe, ":ends")) return rc;
        continue;
      }
    }
    const int jr = ensure_jit(k, fm, 0);
    if (jr < 0) return -jr;
    const int a0 = affs[k].a0, a1 = affs[k].a1;
    const int jb = std::max(a0, fb), je = std::min(a1, fe);
    if (jr > 0 && je > jb) {
      if (int rc = prof_begin(name)) return rc;
      JitArgs a = ja;
      a.begin = jb - a0;
      a.end = je - a0;
      a.out_base = jits[0][k][fm].out_base;
      a.out_stride = jits[0][k][fm].out_stride;
      AB_CUDA(launch_jit(jits[0][k][fm], a, bulk_stream(k)));
      launches++;
      jit_launches++;
      if (int rc = prof_end()) return rc;
      if (int rc = rest(fb, jb, ":ends")) return rc;
      if (int rc = rest(je, fe, ":ends")) return rc;
    } else {
      if (int rc = rest(fb, fe, "")) return rc;
    }
  }
  // Fused kernels.  Only the first and last 64 fused segments share entries with the applications at the mesh ends
  // (evaluated above on the side stream); they are launched after that stream has been joined, so that where both
  // write a value the fused result stands.  Everything between runs concurrently with the mesh ends.
  JitArgs fa;
  long long f_b = 0, f_e = 0;
  // `spread`: the pieces write disjoint rows, so odd pieces go to the second bulk stream and run beside the even ones
  // (the piece holding the shared node leaves room for another block per SM)
  // (measured, sweep s10/s11: with the mesh ends on their own streams the pieces are faster back to back on one stream,
  // 0.558 against 0.597 ms; ASSET_B200_PIECE_OVERLAP=1 brings the second stream back)
  const bool spread_pieces = overlap && getenv("ASSET_B200_PIECE_OVERLAP") && !getenv("ASSET_B200_NO_PIECE_OVERLAP");
  auto fused_launch = [&](long long lo, long long hi, cudaStream_t st, bool spread = false) -> int {
    if (hi <= lo) return 0;
    fa.begin = (int) lo;
    fa.end = (int) hi;
    for (size_t pc = 0; pc < fjits.size(); pc++) {
      const JitKernel& jk = fjits[pc];
      fa.out_base = jk.out_base;
      fa.out_stride = jk.out_stride;
      cudaStream_t use = st;
      if (spread && spread_pieces && (pc & 1)) {
        if (!forked2) {
          if (!forked) cudaEventRecord(ev_fork, stream);
          cudaStreamWaitEvent(bulk2, ev_fork, 0);
          forked2 = true;
        }
        use = bulk2;
      }
      AB_CUDA(launch_jit(jk, fa, use));
      launches++;
      jit_launches++;
    }
    return 0;
  };
  if (fused_eval) {
    if (int rc = prof_begin("fused")) return rc;
    fa.X = d_X;
    fa.L = d_LE;
    fa.kkt = d_kkt;
    fa.fx = d_fxe;
    fa.stage = d_stage;
    fa.appconst = d_appc_all;
    fa.vindex = (const int*) d_LI;  // multipliers of the inequalities (see JitSpec::fin)
    // direct stores address the small output arrays through the map slots (jit.cpp: out_param)
    fa.cindex = (const int*) d_fxi;
    fa.kkt_dest = (const int*) d_pgx;
    fa.f_dest = (const int*) d_agx;
    fa.obj_scale = obj_scale;
    const long long R = fused.t_hi - fused.t_lo;
    f_b = R * rank / world;
    f_e = R * (rank + 1) / world;
    if (windowed) f_b = win_f_b, f_e = win_f_e;
    if (overlap && !forked) {  // fork the side streams before the bulk is queued, or they would wait for it
      cudaEventRecord(ev_fork, stream);
      cudaStreamWaitEvent(side, ev_fork, 0);
      forked = true;
    }
    if (spread_pieces && fjits.size() > 1 && !forked2) {
      cudaStreamWaitEvent(bulk2, ev_fork, 0);
      forked2 = true;
    }
    // the mesh ends write nothing a fused thread owns (redirected maps), so the pieces cover the whole range at once
    if (int rc = fused_launch(f_b, f_e, stream, true)) return rc;
    mark("stream: fused pieces done", stream);
    if (forked2) mark("bulk2: fused pieces done", bulk2);
    if (int rc = prof_end()) return rc;
    pool_gather();
    cudaStream_t es = small_stream();
    mark("side pool: mesh ends done", es);
    // targets that do not depend on the fused kernels are summed here, beside the bulk
    struct {
      DevGather* g;
      double* out;
      bool on;
    } early[3] = {{&gkE, d_kkt, out_kkt}, {&gpE, d_pgx, out_pgx}, {&gaE, d_agx, out_agx}};
    for (auto& eg: early)
      if (eg.on && eg.g->dev[0].ntargets > 0) {
        if (int rc = prof_begin("gather:early")) return rc;
        AB_CUDA(launch_gather(eg.g->dev[0], eg.g->dev[1], d_stage, eg.out, d_gscratch, es));
        launches++;
        if (int rc = prof_end()) return rc;
      }
  }
  if (forked) mark("side: early gathers done", side);
  joined = true;
  if (forked) {
    AB_CUDA(cudaEventRecord(ev_join, side));
    AB_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
  }
  if (forked2) {
    AB_CUDA(cudaEventRecord(ev_join2, bulk2));
    AB_CUDA(cudaStreamWaitEvent(stream, ev_join2, 0));
  }
  if (fused_eval) {
    if (world > 1 && rank == 0 && !windowed) {
      double* arrs[OA_COUNT] = {d_kkt, d_pgx, d_agx, d_fxe, d_fxi, d_val};
      for (int a = 0; a < OA_COUNT; a++)
        if (clear_n[a] > 0) {
          AB_CUDA(launch_clear(arrs[a], d_clear[a], clear_n[a], stream));
          launches++;
        }
    }
  }
  auto G = [&](DevGather& g, double* out, const char* label) -> int {
    const int n = (g.dev[0].ntargets > 0) + (g.dev[1].ntargets > 0);
    if (n == 0) return 0;
    if (int rc = prof_begin(label)) return rc;
    AB_CUDA(launch_gather(g.dev[0], g.dev[1], d_stage, out, d_gscratch, stream));
    launches += n + (g.dev[1].ntargets > 0);
    return prof_end();
  };
  int rc = 0;
  if (out_kkt && (rc = G(fused_eval ? gkF : gk, d_kkt, "gather:KKT"))) return rc;
  if (out_pgx && (rc = G(fused_eval ? gpF : gp, d_pgx, "gather:PGX"))) return rc;
  if (out_agx && (rc = G(fused_eval ? gaF : ga, d_agx, "gather:AGX"))) return rc;
  const bool fsh = fused_eval && (world > 1 || windowed);
  if (out_fx && ((rc = G(fsh ? geF : ge, d_fxe, "gather:FXE")) || (rc = G(fsh ? giF : gi, d_fxi, "gather:FXI")))) return rc;
  if (out_val && (rc = G(fsh ? gvF : gv, d_val, "gather:val"))) return rc;
  if (timeline) {
    mark("end", stream);
    cudaStreamSynchronize(stream);
    for (size_t k = 1; k < marks.size(); k++) {
      float ms = 0;
      cudaEventElapsedTime(&ms, marks[0].second, marks[k].second);
      fprintf(stderr, "[asset_b200 timeline] %-28s %8.3f ms\n", marks[k].first.c_str(), ms);
    }
    for (auto& m: marks) cudaEventDestroy(m.second);
  }
  return 0;
}

// which output arrays an eval mode produces: {KKT, PGX, AGX, FXE, FXI, val}
static void mode_outputs(int mode, bool out[OA_COUNT]) {
  const bool kkt = mode == EM_KKT || mode == EM_KKTNO || mode == EM_SOE || mode == EM_AUG;
  const bool pgx = mode == EM_KKT || mode == EM_AUG || mode == EM_RHS || mode == EM_OGC;
  const bool agx = mode == EM_KKT || mode == EM_KKTNO || mode == EM_AUG || mode == EM_RHS;
  const bool fx = mode != EM_OBJ;
  const bool val = mode != EM_KKTNO && mode != EM_SOE;
  out[OA_KKT] = kkt;
  out[OA_PGX] = pgx;
  out[OA_AGX] = agx;
  out[OA_FXE] = out[OA_FXI] = fx;
  out[OA_VAL] = val;
}

void Plan::finalize_out_ranges() {
  const long long sizes[OA_COUNT] = {prob.nnz(), prob.nvars, prob.nvars, prob.neq, prob.niq, 1};
  for (int a = 0; a < OA_COUNT; a++) {
    out_ranges[a].clear();
    const long long nb = (long long) out_blocks[a].size();
    for (long long q = 0; q < nb;) {
      if (!out_blocks[a][(size_t) q]) {
        q++;
        continue;
      }
      long long e = q;
      while (e < nb && out_blocks[a][(size_t) e]) e++;
      out_ranges[a].push_back({q * OUT_BLOCK, std::min(e * OUT_BLOCK, sizes[a])});
      q = e;
    }
    std::vector<char>().swap(out_blocks[a]);
  }
}

static CutAll make_cut_all(Plan& P) {
  bool on[OA_COUNT];
  mode_outputs(P.last_mode, on);
  double* arrs[OA_COUNT] = {P.d_kkt, P.d_pgx, P.d_agx, P.d_fxe, P.d_fxi, P.d_val};
  CutAll c;
  c.off[0] = 0;
  for (int a = 0; a < OA_COUNT; a++) {
    const long long n = P.last_fused ? P.cutF_n[a] : (long long) P.part.cut_target[a].size();
    c.arr[a] = arrs[a];
    c.target[a] = P.last_fused ? P.d_cutF_target[a] : P.d_cut_target[a];
    c.flag[a] = P.last_fused ? P.d_cutF_flag[a] : P.d_cut_flag[a];
    c.active[a] = on[a] ? 1 : 0;
    c.off[a + 1] = c.off[a] + n;
  }
  c.buf = P.d_cutbuf;
  return c;
}

int Plan::cut_pack() {
  if (world <= 1 || active_cut_count() == 0) return 0;
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(launch_cut_all(make_cut_all(*this), true, stream));
  launches++;
  return 0;
}

int Plan::cut_unpack() {
  if (world <= 1 || active_cut_count() == 0) return 0;
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(launch_cut_all(make_cut_all(*this), false, stream));
  launches++;
  return 0;
}

}  // namespace ab
