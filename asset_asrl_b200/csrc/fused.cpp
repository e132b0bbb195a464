// Owner-writes fused interior (see fused.hpp): template discovery on the analysed problem, then one scalar program for
// everything a thread owns.
#include "fused.hpp"

#include <algorithm>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <tuple>

namespace ab {

namespace {

// One application family of the template with its index maps at the reference thread and their change per thread.
struct InstRef {
  int fi = -1, m = 1;
  long long a_ref = 0;
  bool home = false;
  int reach = 0;  // how many thread blocks past the home block the application's variables extend
  int E = 0;
  std::vector<long long> v0, vst, c0, cst, l0, lst, d0, dst;
};

struct Finder {
  const Problem& P;
  int anchor = -1;
  long long t_ref = 0, s = 0, vbr = 0, kbr = 0, ks = 0;
  std::vector<InstRef> insts;
  std::vector<int> inc;         // incidences per variable over all functions
  std::vector<int> tinc;        // template: incidences of the owned block's variables
  std::vector<long long> iref;  // indptr of the owned block's rows (vs + 1 entries) at t_ref
  std::string why;

  // ASSET_B200_FUSED_BLOCKED=1: accept inputs that advance by a second stride (block-constant control); opt-in until run
  // on a device (written after the round's GPU time was spent; tests/test_fused_flow_model.py checks the whole flow on
  // the CPU, tests/test_zz4_gpu_blocked_fused.py on the device)
  bool satellites = false;

  explicit Finder(const Problem& p) : P(p) {
    if (const char* e = getenv("ASSET_B200_FUSED_BLOCKED")) satellites = std::atoi(e) != 0;
  }

  // stride of every input of application a of function fn per application step; false at the last application
  static bool strides(const FuncRec& fn, long long a, std::vector<long long>& st) {
    if (a + 1 >= fn.nappl) return false;
    st.resize(fn.ir);
    for (int i = 0; i < fn.ir; i++) st[i] = (long long) fn.vindex[(size_t) (a + 1) * fn.ir + i] - fn.vindex[(size_t) a * fn.ir + i];
    return true;
  }

  bool find_template() {
    // anchor: the heaviest function whose applications tile the variables (each owns one contiguous block of its stride);
    // for a phase that is the defect constraint, one application per mesh segment
    std::vector<std::pair<long long, int>> cand;
    for (size_t f = 0; f < P.funcs.size(); f++) {
      const FuncRec& fn = P.funcs[f];
      if (fn.nappl >= 16 && fn.role != 0) cand.emplace_back(-(long long) fn.nappl * fn.triplets_per_app(), (int) f);
    }
    std::sort(cand.begin(), cand.end());
    if (cand.empty()) return fail("no function with enough applications");
    std::vector<long long> st, owned;
    std::string first_why;
    for (const auto& cd: cand) {
      const FuncRec& an = P.funcs[cd.second];
      t_ref = an.nappl / 2;
      std::string w;
      if (!strides(an, t_ref, st)) continue;
      s = 0;
      bool mixed = false;
      for (long long x: st)
        if (x != 0) {
          if (s != 0 && x != s) mixed = true;
          s = x;
        }
      if (mixed && satellites) {
        // Block-constant control (Blocked_ODE_Wrapper.h:8-27, PhaseIndexer.cpp:373-391): the states advance by one stride,
        // the per-segment controls -- stored in a block of their own behind all states -- by another.  The thread owns
        // the states' block; the controls are *satellite* variables: read with their own stride, their rows and gradient
        // entries written like those of the variables no segment owns (directly where an entry has one source, through
        // the staging buffer and the gather lists otherwise).
        long long vlow = -1;
        for (int i = 0; i < an.ir; i++)
          if (st[i] != 0 && (vlow < 0 || an.vindex[(size_t) t_ref * an.ir + i] < vlow)) {
            vlow = an.vindex[(size_t) t_ref * an.ir + i];
            s = st[i];
          }
        mixed = false;
      }
      if (mixed) w = "inputs of " + an.name + " advance by different strides (block-constant control)";
      if (w.empty() && s <= 0) w = an.name + " does not advance through the variables";
      if (w.empty()) {
        // owned inputs: advancing variables that the previous application has not read (v_j(t) == v - s for no j)
        owned.clear();
        for (int i = 0; i < an.ir; i++) {
          if (st[i] != s) continue;
          const long long v = an.vindex[(size_t) t_ref * an.ir + i];
          bool again = false;
          for (int j = 0; j < an.ir; j++)
            if (st[j] == s && an.vindex[(size_t) t_ref * an.ir + j] == v - s) again = true;
          if (!again) owned.push_back(v);
        }
        std::sort(owned.begin(), owned.end());
        owned.erase(std::unique(owned.begin(), owned.end()), owned.end());
        if ((long long) owned.size() != s || owned.back() - owned.front() + 1 != s)
          w = "the variables an application of " + an.name + " owns are not one contiguous block of its stride";
      }
      if (w.empty()) {
        anchor = cd.second;
        break;
      }
      if (first_why.empty()) first_why = w;
    }
    if (anchor < 0) return fail(first_why.empty() ? "no anchor function" : first_why);
    vbr = owned.front();
    if (vbr + s > P.nvars) return fail("owned block out of range");
    // incidences of every variable and the applications touching the reference block
    inc.assign((size_t) P.nvars, 0);
    std::vector<std::pair<int, long long>> hits;
    for (size_t f = 0; f < P.funcs.size(); f++) {
      const FuncRec& fn = P.funcs[f];
      for (long long a = 0; a < fn.nappl; a++) {
        bool hit = false;
        for (int i = 0; i < fn.ir; i++) {
          const int v = fn.vindex[(size_t) a * fn.ir + i];
          inc[(size_t) v]++;
          if (v >= vbr && v < vbr + s) hit = true;
        }
        if (hit) hits.emplace_back((int) f, a);
      }
    }
    tinc.assign((size_t) s, 0);
    for (long long j = 0; j < s; j++) tinc[(size_t) j] = inc[(size_t) (vbr + j)];
    iref.resize((size_t) s + 1);
    for (long long j = 0; j <= s; j++) iref[(size_t) j] = P.indptr[(size_t) (vbr + j)];
    kbr = iref[0];
    ks = iref[(size_t) s] - iref[0];
    if (vbr + 2 * s > P.nvars || P.indptr[(size_t) (vbr + 2 * s)] - P.indptr[(size_t) (vbr + s)] != ks)
      return fail("CSR rows of consecutive segments differ in length");
    for (auto& h: hits) {
      const FuncRec& fn = P.funcs[h.first];
      InstRef q;
      q.fi = h.first;
      q.a_ref = h.second;
      std::vector<long long> fs;
      if (!strides(fn, q.a_ref, fs)) return fail("a boundary application touches the interior block (" + fn.name + ")");
      long long sf = 0;
      if (satellites) {
        // the stride of the lowest advancing variable (a state) counts; check() verifies every thread afterwards
        long long vlow = -1;
        for (int i = 0; i < fn.ir; i++)
          if (fs[i] != 0 && (vlow < 0 || fn.vindex[(size_t) q.a_ref * fn.ir + i] < vlow)) {
            vlow = fn.vindex[(size_t) q.a_ref * fn.ir + i];
            sf = fs[i];
          }
      } else {
        for (long long x: fs)
          if (x != 0) {
            if (sf != 0 && x != sf) return fail("inputs of " + fn.name + " advance by different strides");
            sf = x;
          }
      }
      if (sf <= 0 || s % sf != 0) return fail("stride of " + fn.name + " does not divide the segment stride");
      q.m = (int) (s / sf);
      if (q.a_ref + q.m >= fn.nappl) return fail("template application at the end of " + fn.name);
      q.E = fn.triplets_per_app();
      const long long a1 = q.a_ref + q.m;
      auto grab = [&](const int* base, int w, std::vector<long long>& x0, std::vector<long long>& xs) {
        x0.resize(w);
        xs.resize(w);
        for (int i = 0; i < w; i++) {
          x0[i] = base[(size_t) q.a_ref * w + i];
          xs[i] = (long long) base[(size_t) a1 * w + i] - x0[i];
        }
      };
      grab(fn.vindex.data(), fn.ir, q.v0, q.vst);
      if (fn.role != 0) grab(fn.cindex.data(), fn.orows, q.c0, q.cst);
      grab(P.kktloc.data() + P.trip_base[h.first], q.E, q.l0, q.lst);
      grab(P.kkt_dest.data() + P.trip_base[h.first], q.E, q.d0, q.dst);
      long long vmin = -1, vmax = -1;
      for (int i = 0; i < fn.ir; i++)
        if (q.vst[i] != 0) {
          if (q.vst[i] != s) {
            if (!satellites) return fail("input of " + fn.name + " does not follow the segment stride");
            continue;  // a satellite variable: no part in where the application lives
          }
          vmin = vmin < 0 ? q.v0[i] : std::min(vmin, q.v0[i]);
          vmax = std::max(vmax, q.v0[i]);
        }
      q.home = vmin >= vbr && vmin < vbr + s;
      q.reach = (int) ((vmax - vbr) / s);
      if (vmax < vbr) q.reach = 0;
      if (q.home && fn.role != 0) {
        if (!fn.unique_constraints) return fail(fn.name + " accumulates into shared constraint rows");
        for (int k = 0; k < fn.orows; k++)
          if (P.f_dest[(size_t) (P.fdest_base[h.first] + q.a_ref * fn.orows + k)] != fn.cindex[(size_t) q.a_ref * fn.orows + k])
            return fail(fn.name + " has staged constraint rows");
      }
      insts.push_back(std::move(q));
    }
    return true;
  }

  bool fail(const std::string& w) {
    why = w;
    return false;
  }

  // does thread t (anchor application t) have exactly the template's structure?
  bool check(long long t) const {
    const long long dt = t - t_ref;
    const long long vb_t = vbr + dt * s;
    if (vb_t < 0 || vb_t + s > P.nvars) return false;
    for (long long j = 0; j < s; j++)
      if (inc[(size_t) (vb_t + j)] != tinc[(size_t) j]) return false;
    for (long long j = 0; j <= s; j++)
      if ((long long) P.indptr[(size_t) (vb_t + j)] != iref[(size_t) j] + dt * ks) return false;
    const long long SV = P.niq;
    for (long long j = 0; j < s; j++)  // the solver's primal diagonal entry is the first slot of the row
      if (P.kktloc[(size_t) (P.n_user + SV + vb_t + j)] != P.indptr[(size_t) (vb_t + j)]) return false;
    for (const InstRef& q: insts) {
      const FuncRec& fn = P.funcs[q.fi];
      const long long a = q.a_ref + q.m * dt;
      if (a < 0 || a >= fn.nappl) return false;
      const int* V = fn.vindex.data() + (size_t) a * fn.ir;
      for (int i = 0; i < fn.ir; i++)
        if (V[i] != q.v0[i] + dt * q.vst[i]) return false;
      if (fn.role != 0) {
        const int* Cx = fn.cindex.data() + (size_t) a * fn.orows;
        for (int k = 0; k < fn.orows; k++) {
          if (Cx[k] != q.c0[k] + dt * q.cst[k]) return false;
          if (q.home && P.f_dest[(size_t) (P.fdest_base[q.fi] + a * fn.orows + k)] != Cx[k]) return false;
        }
      }
      const int* L = P.kktloc.data() + P.trip_base[q.fi] + a * q.E;
      const int* D = P.kkt_dest.data() + P.trip_base[q.fi] + a * q.E;
      for (int e = 0; e < q.E; e++) {
        if (L[e] != q.l0[e] + dt * q.lst[e]) return false;
        if (D[e] != q.d0[e] + dt * q.dst[e]) return false;
      }
    }
    return true;
  }
};

}  // namespace

}  // namespace ab

namespace ab {

namespace {

// Copies the sub-DAG of `src` reachable from `roots` into `dst`, leaves replaced through `leaf`.
struct DagCopy {
  const SDag& src;
  SDag& dst;
  std::vector<int> map;
  template<class LeafFn>
  DagCopy(const SDag& s, SDag& d, const std::vector<int>& roots, LeafFn&& leaf) : src(s), dst(d), map(s.n.size(), -1) {
    std::vector<char> need(s.n.size(), 0);
    for (int r: roots)
      if (r >= 0) need[(size_t) r] = 1;
    for (long long i = (long long) s.n.size() - 1; i >= 0; i--) {  // operands have smaller ids than their users
      if (!need[(size_t) i]) continue;
      const SNode& nd = s.n[(size_t) i];
      if (nd.op == OP_CONST || nd.op == OP_INPUT || nd.op == OP_LAMBDA || nd.op == OP_APPCONST) continue;
      for (int o: {nd.a, nd.b, nd.c})
        if (o >= 0) need[(size_t) o] = 1;
    }
    for (size_t i = 0; i < s.n.size(); i++) {
      if (!need[i]) continue;
      const SNode& nd = s.n[i];
      auto m = [&](int o) { return o >= 0 ? map[(size_t) o] : -1; };
      switch (nd.op) {
        case OP_CONST: map[i] = d.constant(nd.imm); break;
        case OP_INPUT:
        case OP_LAMBDA:
        case OP_APPCONST: map[i] = leaf(nd.op, nd.a); break;
        case OP_ADD: map[i] = d.add(m(nd.a), m(nd.b)); break;
        case OP_SUB: map[i] = d.sub(m(nd.a), m(nd.b)); break;
        case OP_MUL: map[i] = d.mul(m(nd.a), m(nd.b)); break;
        case OP_DIV: map[i] = d.div(m(nd.a), m(nd.b)); break;
        case OP_NEG: map[i] = d.neg(m(nd.a)); break;
        case OP_ADDC: map[i] = d.addc(m(nd.a), nd.imm); break;
        case OP_MULC: map[i] = d.mulc(m(nd.a), nd.imm); break;
        case OP_TABIDX:
        case OP_TABLOAD: map[i] = d.raw(nd.op, m(nd.a), -1, -1, d.tab_import(s, nd.imm)); break;  // table data move with the node
        default: map[i] = d.raw(nd.op, m(nd.a), m(nd.b), m(nd.c), nd.imm); break;
      }
    }
  }
  int operator()(int node) const { return map.at((size_t) node); }
};

}  // namespace

FusedPlan build_fused(const Problem& prob, int min_threads, int npieces) {
  FusedPlan FP;
  if (npieces == 0)
    if (const char* e = getenv("ASSET_B200_FUSED_PIECES")) npieces = std::max(0, std::atoi(e));
  if (!prob.analyzed) {
    FP.why = "problem not analysed";
    return FP;
  }
  Finder F(prob);
  if (!F.find_template()) {
    FP.why = F.why;
    return FP;
  }
  if (!F.check(F.t_ref) || !F.check(F.t_ref + 1)) {
    FP.why = "the mesh is not periodic around its middle";
    return FP;
  }
  long long lo = F.t_ref, hi = F.t_ref + 1;
  while (F.check(lo - 1)) lo--;
  while (F.check(hi)) hi++;
  if (hi - lo < min_threads) {
    FP.why = "periodic interior of " + std::to_string(hi - lo) + " segments is too short";
    return FP;
  }
  const long long s = F.s, dlo = lo - F.t_ref;
  FP.anchor = F.anchor;
  FP.t_lo = (int) lo;
  FP.t_hi = (int) hi;
  FP.vs = s;
  FP.vb = F.vbr + dlo * s;
  FP.ks = F.ks;
  FP.kb = F.kbr + dlo * F.ks;

  // ---- applications the per-function path no longer evaluates
  const size_t nf = prob.funcs.size();
  FP.ex_lo.assign(nf, 0);
  FP.ex_hi.assign(nf, 0);
  {
    std::vector<long long> xlo(nf, -1), xhi(nf, -1), cnt(nf, 0);
    std::vector<int> mm(nf, 0);
    for (const InstRef& q: F.insts) {
      if (!q.home) continue;
      const long long T = (hi - q.reach) - lo;  // threads whose application of this family stays inside the fused rows
      if (T <= 0) continue;
      const long long a_first = q.a_ref + q.m * dlo, a_last = q.a_ref + q.m * (hi - q.reach - 1 - F.t_ref);
      xlo[q.fi] = xlo[q.fi] < 0 ? a_first : std::min(xlo[q.fi], a_first);
      xhi[q.fi] = std::max(xhi[q.fi], a_last + 1);
      cnt[q.fi] += T;
      mm[q.fi] = q.m;
    }
    for (size_t f = 0; f < nf; f++) {
      if (xlo[f] < 0) continue;
      if (xhi[f] - xlo[f] != cnt[f]) {
        FP.why = "home applications of " + prob.funcs[f].name + " do not tile a contiguous range";
        return FP;
      }
      FP.ex_lo[f] = (int) xlo[f];
      FP.ex_hi[f] = (int) xhi[f];
    }
  }
  // per-application constants of all functions, concatenated
  FP.appc_offset.assign(nf, 0);
  {
    long long off = 0;
    for (size_t f = 0; f < nf; f++) {
      FP.appc_offset[f] = off;
      off += (long long) prob.funcs[f].appconst.size();
    }
  }

  // ---- the program
  ScalarFn fused;
  SDag& D = fused.dag;
  std::map<std::tuple<int, long long, long long>, int> fin_index;
  auto input = [&](int arr, long long base_ref, long long stride) {
    const long long base = base_ref + dlo * stride;  // rebased to rel = 0
    auto key = std::make_tuple(arr, base, stride);
    auto it = fin_index.find(key);
    if (it == fin_index.end()) {
      FP.fin.push_back({arr, base, stride});
      it = fin_index.emplace(key, (int) FP.fin.size() - 1).first;
    }
    return D.input(it->second);
  };
  const long long ks = F.ks;
  std::vector<std::vector<int>> kk((size_t) ks), ag((size_t) s), pg((size_t) s);
  // what each contribution to a matrix entry is (parallel to kk): the partial-matrix modes sum a subset
  enum : unsigned char { T_OBJ = 1, T_HESS = 2, T_SOLVER = 4 };
  std::vector<std::vector<unsigned char>> kk_tag((size_t) ks);
  struct Extra {
    JitOut where;
    int node;
    int kind = 0;  // far_out: 0 matrix entry, 1 gradient entry, 2 objective value
    int role = 1;
    unsigned char tag = 0;  // kind 0: T_OBJ / T_HESS
  };
  std::vector<Extra> fx_out, far_out;
  const long long SV = prob.niq;
  for (long long j = 0; j < s; j++) {  // solver-owned primal diagonal: first slot of the row
    kk[(size_t) (F.iref[(size_t) j] - F.kbr)].push_back(input(FA_STAGE, prob.stage_solver + SV + F.vbr + j, s));
    kk_tag[(size_t) (F.iref[(size_t) j] - F.kbr)].push_back(T_SOLVER);
  }

  std::map<int, ScalarFn> work;  // per function: lowered function with first and second derivatives
  for (const InstRef& q: F.insts) {
    const FuncRec& fn = prob.funcs[q.fi];
    auto wit = work.find(q.fi);
    if (wit == work.end()) {
      const Tape& tp = prob.tapes.at(fn.tape);
      if (!tp.fn) {
        FP.why = "no lowered function for " + fn.name;
        return FP;
      }
      ScalarFn w = *tp.fn;
      if ((int) w.grads.size() != w.nin) w.make_gradient();
      if (w.jac.empty()) w.make_second_order();
      wit = work.emplace(q.fi, std::move(w)).first;
    }
    const ScalarFn& cf = wit->second;
    const int ir = fn.ir, orr = fn.orows;
    const bool lay_h = !fn.linear, lay_j = fn.role != 0;
    std::vector<int> roots;
    roots.insert(roots.end(), cf.outs.begin(), cf.outs.end());
    roots.insert(roots.end(), cf.grads.begin(), cf.grads.end());
    if (lay_j) roots.insert(roots.end(), cf.jac.begin(), cf.jac.end());
    if (lay_h) roots.insert(roots.end(), cf.hess.begin(), cf.hess.end());
    DagCopy cp(cf.dag, D, roots, [&](uint16_t op, int k) {
      if (op == OP_INPUT) return input(FA_X, q.v0.at(k), q.vst.at(k));
      if (op == OP_LAMBDA) {
        if (fn.role == 0) return input(FA_SCALE, 0, 0);
        return input(fn.role == 1 ? FA_LE : FA_LI, q.c0.at(k), q.cst.at(k));
      }
      return input(FA_APPC, FP.appc_offset[q.fi] + q.a_ref * fn.nappconst + k, (long long) q.m * fn.nappconst);
    });
    auto owned_var = [&](int i) { return q.vst[i] == s && q.v0[i] >= F.vbr && q.v0[i] < F.vbr + s; };
    auto place = [&](int e, int row_in, int node, bool hess) {
      const unsigned char tag = (unsigned char) ((fn.role == 0 ? T_OBJ : 0) | (hess ? T_HESS : 0));
      if (q.lst[e] == ks && q.l0[e] >= F.kbr && q.l0[e] < F.kbr + ks) {
        kk[(size_t) (q.l0[e] - F.kbr)].push_back(node);
        kk_tag[(size_t) (q.l0[e] - F.kbr)].push_back(tag);
      } else if (q.home && q.vst[row_in] != s) {
        // row of a variable no segment owns (parameters, the phase's end times; with ASSET_B200_FUSED_BLOCKED the
        // per-segment controls, whose stride differs): as in the per-function kernels
        if (q.d0[e] >= 0)
          far_out.push_back({{FA_KKT, q.d0[e] + dlo * q.dst[e], q.dst[e]}, node, 0, fn.role, tag});
        else
          far_out.push_back({{FA_STAGE, ~q.d0[e] - dlo * q.dst[e], -q.dst[e]}, node, 0, fn.role, tag});
      }
    };
    int e = 0;
    for (int i = 0; i < ir; i++) {
      if (lay_h)
        for (int j = i; j < ir; j++, e++) place(e, q.v0[i] <= q.v0[j] ? i : j, cp(cf.hess[(size_t) j * ir + i]), true);
      if (lay_j)
        for (int k = 0; k < orr; k++, e++) place(e, i, cp(cf.jac[(size_t) k * ir + i]), false);
      const int g = cp(cf.grads[(size_t) i]);
      if (owned_var(i)) {
        (fn.role == 0 ? pg : ag)[(size_t) (q.v0[i] - F.vbr)].push_back(g);
      } else if (q.home && q.vst[i] != s) {
        far_out.push_back({{FA_STAGE, prob.g_stage[q.fi] + (q.a_ref + q.m * dlo) * ir + i, (long long) q.m * ir}, g, 1, fn.role});
      }
    }
    if (q.home) {
      FP.n_inst_home++;
      for (int k = 0; k < orr; k++) {
        const int v = cp(cf.outs[(size_t) k]);
        if (fn.role == 0)
          far_out.push_back({{FA_STAGE, prob.v_stage[q.fi] + q.a_ref + q.m * dlo, q.m}, D.mul(v, input(FA_SCALE, 0, 0)), 2, 0});
        else
          fx_out.push_back({{fn.role == 1 ? FA_FXE : FA_FXI, q.c0[k] + dlo * q.cst[k], q.cst[k]}, v});
      }
    }
    FP.insts.push_back({q.fi, q.m, q.a_ref + q.m * dlo, q.home});
  }

  // ---- outputs in address order: the thread's CSR region front to back, then the small arrays; cut into row blocks
  auto sum = [&](const std::vector<int>& v) {
    int acc = v[0];
    for (size_t k = 1; k < v.size(); k++) acc = D.add(acc, v[k]);
    return acc;
  };
  struct Rec {
    JitOut where;
    int node;
    long long row;  // owned row the output belongs to (-1: none, goes with the last piece)
  };
  std::vector<Rec> recs;
  for (long long o = 0; o < ks; o++) {
    if (kk[(size_t) o].empty()) {
      FP.why = "a CSR slot of the owned rows has no source";
      return FP;
    }
    const long long row = (std::upper_bound(F.iref.begin(), F.iref.end(), F.kbr + o) - F.iref.begin()) - 1;
    recs.push_back({{FA_KKT, FP.kb + o, ks}, sum(kk[(size_t) o]), row});
  }
  std::sort(fx_out.begin(), fx_out.end(), [](const Extra& a, const Extra& b) {
    return std::make_tuple(a.where.arr, a.where.base) < std::make_tuple(b.where.arr, b.where.base);
  });
  for (const Extra& x: fx_out) recs.push_back({x.where, x.node, -1});
  for (long long j = 0; j < s; j++)
    if (!ag[(size_t) j].empty()) recs.push_back({{FA_AGX, FP.vb + j, s}, sum(ag[(size_t) j]), j});
  for (long long j = 0; j < s; j++)
    if (!pg[(size_t) j].empty()) recs.push_back({{FA_PGX, FP.vb + j, s}, sum(pg[(size_t) j]), j});
  for (const Extra& x: far_out) recs.push_back({x.where, x.node, -1});
  fused.nin = (int) FP.fin.size();
  fused.nout = (int) recs.size();
  // pieces are blocks of whole owned rows; bnd holds the np+1 row boundaries
  std::vector<long long> final_bnd;  // the row split the evalKKT pieces ended up with
  auto build_pieces = [&](const std::vector<long long>& bnd) {
    final_bnd = bnd;
    const int np = (int) bnd.size() - 1;
    FP.pieces.assign((size_t) np, FusedPlan::Piece());
    std::vector<std::vector<ProgOut>> outs((size_t) np);
    for (const Rec& r: recs) {
      const int pc = r.row < 0 ? np - 1 : (int) (std::upper_bound(bnd.begin() + 1, bnd.end() - 1, r.row) - bnd.begin()) - 1;
      FP.pieces[(size_t) pc].fout.push_back(r.where);
      outs[(size_t) pc].push_back({r.node, PO_TABLE, (uint32_t) FP.pieces[(size_t) pc].fout.size() - 1});
    }
    for (int pc = 0; pc < np; pc++) FP.pieces[(size_t) pc].prog = emit_program_outs(fused, outs[(size_t) pc]);
  };
  auto even = [&](int np) {
    std::vector<long long> bnd((size_t) np + 1);
    for (int i = 0; i <= np; i++) bnd[(size_t) i] = (s * i + np - 1) / np;
    return bnd;
  };
  auto cost = [&](int pc) { return (double) FP.pieces[(size_t) pc].prog.n_fma + 2.0 * (double) FP.pieces[(size_t) pc].fout.size(); };
  // Experimental (ASSET_B200_FUSED_BALANCE=1): the rows are far from equally expensive (the first node's rows of an
  // LGL5 segment carry twice the arithmetic of the second's); move the boundaries towards equal-cost quantiles, taking
  // the cost of a piece as spread evenly over its rows, and keep the split whose largest piece is smallest.  Off by
  // default: the pieces run back to back at full width (sweep r2h: two-stream overlap gains 2%), so their sum counts,
  // not the largest, and the even split has the smaller sum.
  auto balance = [&](int np) {
    std::vector<long long> bnd = even(np), best = bnd;
    double best_max = 1e300;
    for (int it = 0; it < 4; it++) {
      build_pieces(bnd);
      double total = 0, mx = 0;
      for (int pc = 0; pc < np; pc++) total += cost(pc), mx = std::max(mx, cost(pc));
      if (mx < best_max) best_max = mx, best = bnd;
      std::vector<double> cum((size_t) s + 1, 0.0);  // cumulative cost at each row boundary
      for (int pc = 0; pc < np; pc++) {
        const long long lo = bnd[(size_t) pc], hi = bnd[(size_t) pc + 1];
        for (long long r = lo; r < hi; r++) cum[(size_t) r + 1] = cum[(size_t) r] + cost(pc) / (double) (hi - lo);
      }
      std::vector<long long> nb((size_t) np + 1, 0);
      nb[(size_t) np] = s;
      for (int i = 1; i < np; i++) {
        const double want = total * i / np;
        long long r = std::lower_bound(cum.begin(), cum.end(), want) - cum.begin();
        if (r > 0 && want - cum[(size_t) r - 1] < cum[(size_t) r] - want) r--;
        nb[(size_t) i] = std::min<long long>(s - (np - i), std::max<long long>(nb[(size_t) i - 1] + 1, r));
      }
      if (nb == bnd) break;
      bnd = nb;
    }
    if (best != bnd || FP.pieces.size() != (size_t) np) build_pieces(best);
    else {
      double mx = 0;
      for (int pc = 0; pc < np; pc++) mx = std::max(mx, cost(pc));
      if (mx > best_max) build_pieces(best);
    }
  };
  // Greedy split by on-chip state: a piece takes as many whole rows as keep its parked values within `slot_limit`
  // lane-private shared-memory slots (256 = 64 KB per 32-thread block, three blocks per SM), so that the number of
  // pieces -- each re-reads the inputs and recomputes what it shares with its neighbours -- is the smallest the shared
  // memory allows.  Sweeps s10 / s16: LGL5 low-thrust is fastest as two pieces although four would fit four blocks per
  // SM; LGL7 wants its heavy first rows (the node shared with the previous segment) in pieces of their own.
  auto slots_of = [&](long long r0, long long r1) {
    std::vector<ProgOut> outs;
    FusedPlan::Piece pc;
    for (const Rec& r: recs) {
      const bool in = r.row < 0 ? r1 == s : (r.row >= r0 && r.row < r1);
      if (!in) continue;
      pc.fout.push_back(r.where);
      outs.push_back({r.node, PO_TABLE, (uint32_t) pc.fout.size() - 1});
    }
    pc.prog = emit_program_outs(fused, outs);
    FusedPlan tmp;
    tmp.fin = FP.fin;
    tmp.t_lo = FP.t_lo;
    tmp.t_hi = FP.t_hi;
    tmp.pieces.push_back(std::move(pc));
    JitStats st;
    fused_ptx(tmp, 0, nullptr, &st);
    return st.parked_slots;
  };
  auto by_slots = [&](int slot_limit) {
    std::vector<long long> bnd {0};
    while (bnd.back() < s) {
      const long long r0 = bnd.back();
      long long lo = r0 + 1, hi = s;  // the largest r1 in [lo, hi] that fits (lo always taken)
      if (slots_of(r0, s) <= slot_limit) lo = s;
      while (lo < hi) {
        const long long mid = (lo + hi + 1) / 2;
        if (slots_of(r0, mid) <= slot_limit) lo = mid;
        else hi = mid - 1;
      }
      bnd.push_back(lo);
      if (bnd.size() > 17) break;
    }
    if (bnd.back() != s) bnd.push_back(s);
    return bnd;
  };
  try {
    if (npieces == 0 && !getenv("ASSET_B200_FUSED_EVEN_SPLIT")) {
      int limit = 256;
      if (const char* e = getenv("ASSET_B200_FUSED_SLOTS")) limit = std::max(16, std::atoi(e));
      build_pieces(by_slots(limit));
    } else if (npieces > 0) {
      if (getenv("ASSET_B200_FUSED_BALANCE")) balance((int) std::min<long long>(npieces, s));
      else build_pieces(even((int) std::min<long long>(npieces, s)));
    } else {
      // one piece per ~1300 FMAs of the whole program (sweep r2g, 32-thread blocks with zero runs written out: four
      // pieces for the 5090-FMA LGL5 low-thrust segment run the step in 0.81 ms, two in 0.88, one leaves a single
      // block per SM); at most six (LGL7 low-thrust, 100 k segments: 1.17 ms with six, 1.20 with eight, 1.22 with four;
      // the small programs -- LGL3, trapezoidal, VanDerPol, HyperSens -- are fastest as one piece; sweep r2q)
      build_pieces(even(1));
      const int np = std::min<long long>(std::min<long long>(6, s), std::max(1, (FP.pieces[0].prog.n_fma + 650) / 1300));
      if (np > 1) {
        if (getenv("ASSET_B200_FUSED_BALANCE")) balance(np);
        else build_pieces(even(np));
      }
    }
  } catch (const std::exception& ex) {
    FP.why = std::string("fused program: ") + ex.what();
    return FP;
  }
  // residual-only programs (no matrix entries): one piece each
  try {
    for (int mode = 4; mode <= 7; mode++) {
      const bool want_ag = mode == 4, want_pg = mode == 4 || mode == 5, want_fx = mode != 7;
      FusedPlan::Piece pc;
      std::vector<ProgOut> outs;
      auto add = [&](const JitOut& where, int node) {
        pc.fout.push_back(where);
        outs.push_back({node, PO_TABLE, (uint32_t) pc.fout.size() - 1});
      };
      if (want_fx)
        for (const Extra& x: fx_out) add(x.where, x.node);
      if (want_ag)
        for (long long j = 0; j < s; j++)
          if (!ag[(size_t) j].empty()) add({FA_AGX, FP.vb + j, s}, sum(ag[(size_t) j]));
      if (want_pg)
        for (long long j = 0; j < s; j++)
          if (!pg[(size_t) j].empty()) add({FA_PGX, FP.vb + j, s}, sum(pg[(size_t) j]));
      for (const Extra& x: far_out) {
        if (x.kind == 0) continue;
        if (x.kind == 1 && !(x.role == 0 ? want_pg : want_ag)) continue;
        add(x.where, x.node);
      }
      if (outs.empty()) continue;
      pc.prog = emit_program_outs(fused, outs);
      FP.residual[mode].push_back(std::move(pc));
    }
  } catch (const std::exception& ex) {
    for (auto& r: FP.residual) r.clear();  // the per-function kernels evaluate these modes
  }
  // Partial-matrix evaluations on the same threads and the same row blocks (evalKKTNO 1: everything but the objective;
  // evalSOE 2: constraint values and Jacobians; evalAUG 3: Jacobians, both gradients, values -- NonLinearProgram.cpp:
  // 540-666): every owned matrix entry is written, with the sum of the contributions the mode computes (the solver's
  // diagonal coefficient always; zero if nothing is left).
  try {
    const int np = (int) final_bnd.size() - 1;
    for (int mode = 1; mode <= 3 && np >= 1; mode++) {
      const bool want_ag = mode == 1 || mode == 3, want_pg = mode == 3, want_val = mode == 3;
      auto allowed = [&](unsigned char tag) {
        if (tag & T_SOLVER) return true;
        if (mode == 1) return !(tag & T_OBJ);
        return !(tag & T_HESS) && !(tag & T_OBJ);  // Jacobians of the constraints only
      };
      std::vector<Rec> mrecs;
      for (long long o = 0; o < ks; o++) {
        int acc = -1;
        for (size_t k = 0; k < kk[(size_t) o].size(); k++) {
          if (!allowed(kk_tag[(size_t) o][k])) continue;
          acc = acc < 0 ? kk[(size_t) o][k] : D.add(acc, kk[(size_t) o][k]);
        }
        if (acc < 0) acc = D.constant(0.0);
        const long long row = (std::upper_bound(F.iref.begin(), F.iref.end(), F.kbr + o) - F.iref.begin()) - 1;
        mrecs.push_back({{FA_KKT, FP.kb + o, ks}, acc, row});
      }
      for (const Extra& x: fx_out) mrecs.push_back({x.where, x.node, -1});
      if (want_ag)
        for (long long j = 0; j < s; j++)
          if (!ag[(size_t) j].empty()) mrecs.push_back({{FA_AGX, FP.vb + j, s}, sum(ag[(size_t) j]), j});
      if (want_pg)
        for (long long j = 0; j < s; j++)
          if (!pg[(size_t) j].empty()) mrecs.push_back({{FA_PGX, FP.vb + j, s}, sum(pg[(size_t) j]), j});
      for (const Extra& x: far_out) {
        if (x.kind == 0 && !allowed(x.tag)) continue;
        if (x.kind == 1 && !(x.role == 0 ? want_pg : want_ag)) continue;
        if (x.kind == 2 && !want_val) continue;
        mrecs.push_back({x.where, x.node, -1});
      }
      std::vector<FusedPlan::Piece> pcs((size_t) np);
      std::vector<std::vector<ProgOut>> outs((size_t) np);
      for (const Rec& r: mrecs) {
        const int pc = r.row < 0 ? np - 1 : (int) (std::upper_bound(final_bnd.begin() + 1, final_bnd.end() - 1, r.row) - final_bnd.begin()) - 1;
        pcs[(size_t) pc].fout.push_back(r.where);
        outs[(size_t) pc].push_back({r.node, PO_TABLE, (uint32_t) pcs[(size_t) pc].fout.size() - 1});
      }
      for (int pc = 0; pc < np; pc++) pcs[(size_t) pc].prog = emit_program_outs(fused, outs[(size_t) pc]);
      FP.residual[mode] = std::move(pcs);
    }
  } catch (const std::exception& ex) {
    for (int mode = 1; mode <= 3; mode++) FP.residual[mode].clear();
  }
  FP.ok = true;
  return FP;
}

std::string fused_ptx(const FusedPlan& fp, int piece, std::vector<JitOut>* outs, JitStats* stats, int reg_budget, int mode) {
  JitSpec sp;
  const std::vector<FusedPlan::Piece>& pieces = mode == 0 ? fp.pieces : fp.residual[mode & 7];
  sp.name = mode == 0 ? "ab_fused_kkt" + std::to_string(piece) : "ab_fused_m" + std::to_string(mode) + "_" + std::to_string(piece);
  sp.role = 1;  // multipliers and the objective scale are table entries, not the role's implicit operand
  sp.indexed = false;
  // One warp per block and every zero between stored values written (ncu, capture r2f): with the structural zeros left
  // out the region has holes, and partly written sectors cost 1.03 GB of DRAM fill reads per 200 k segments; written
  // out, the region is one contiguous stream (reads drop to 0.25 GB) and the kernels become latency-bound, where the
  // smaller block's third resident warp per SM helps.
  sp.nthreads = 32;
  sp.zero_run_min = 1000;
  // outputs leave straight from registers as whole 32-byte sectors (jit.hpp: direct_stores); ASSET_B200_FUSED_DIRECT=0
  // brings back the shared-memory transposition walk for comparison
  sp.direct_stores = true;
  if (const char* e = getenv("ASSET_B200_FUSED_DIRECT")) sp.direct_stores = std::atoi(e) != 0;
  sp.fin = &fp.fin;
  sp.fout = &pieces.at((size_t) piece).fout;
  jit_env_knobs(sp);
  // knobs of the fused kernels alone (sweeps)
  if (const char* e = getenv("ASSET_B200_FUSED_ZRUN")) sp.zero_run_min = std::max(1, std::atoi(e));
  if (const char* e = getenv("ASSET_B200_FUSED_REGS")) sp.reg_budget = std::max(16, std::atoi(e));
  if (const char* e = getenv("ASSET_B200_FUSED_REGS_BY_PIECE")) {  // comma-separated register budgets, one per piece
    std::string list(e);
    size_t pos = 0;
    for (int k = 0; pos <= list.size(); k++) {
      const size_t end = std::min(list.find(',', pos), list.size());
      if (k == piece && end > pos) sp.reg_budget = std::max(16, std::atoi(list.substr(pos, end - pos).c_str()));
      pos = end + 1;
    }
  }
  if (const char* e = getenv("ASSET_B200_FUSED_THREADS")) sp.nthreads = std::max(32, std::atoi(e) / 32 * 32);
  if (reg_budget > 0) sp.reg_budget = reg_budget;
  if (stats) stats->reg_budget = sp.reg_budget;
  AffineMaps am;
  am.a0 = 0;
  am.a1 = fp.t_hi - fp.t_lo;
  return emit_ptx(pieces.at((size_t) piece).prog, sp, am, outs, stats);
}

}  // namespace ab
