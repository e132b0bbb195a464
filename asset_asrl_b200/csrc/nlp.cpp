#include "nlp.hpp"

#include <algorithm>
#include <stdexcept>
#include <thread>

namespace ab {

namespace {

// Enumerate the KKT triplets of one function in the reference's emission order
// (DenseFunctionBase.h:1097-1129): per application, per input column i: lower-triangular Hessian
// rows j>=i (non-linear functions only), then Jacobian rows (constraints only).
template<class F>
inline void for_each_triplet(const FuncRec& fn, int conoffset, F&& cb) {
  const bool hess = !fn.linear;
  const bool jac = fn.role != 0;
  const int ir = fn.ir, orr = fn.orows;
  for (int V = 0; V < fn.nappl; V++) {
    const int* vi = fn.vindex.data() + (size_t) V * ir;
    const int* ci = fn.cindex.data() + (size_t) V * orr;
    for (int i = 0; i < ir; i++) {
      if (hess)
        for (int j = i; j < ir; j++) cb(vi[j], vi[i]);
      if (jac)
        for (int j = 0; j < orr; j++) cb(ci[j] + conoffset, vi[i]);
    }
  }
}

void build_gather(const std::vector<int>& tgt_of_src,  // target id per source (in ascending stage order)
                  const std::vector<int>& stage_of_src, int ntargets_total,
                  const std::vector<int>& out_index,    // output index per target id
                  const std::vector<char>& keep,        // which target ids get a gather entry
                  GatherList out[2]) {
  std::vector<long long> cnt((size_t) ntargets_total + 1, 0);
  for (int t: tgt_of_src) cnt[(size_t) t + 1]++;
  for (int t = 0; t < ntargets_total; t++) cnt[(size_t) t + 1] += cnt[t];
  std::vector<int> sorted(tgt_of_src.size());
  {
    std::vector<long long> pos(cnt.begin(), cnt.end() - 1);
    for (size_t s = 0; s < tgt_of_src.size(); s++) sorted[(size_t) pos[tgt_of_src[s]]++] = stage_of_src[s];
  }
  for (int w = 0; w < 2; w++) {
    out[w].ptr.assign(1, 0);
    out[w].src.clear();
    out[w].target.clear();
  }
  for (int t = 0; t < ntargets_total; t++) {
    if (!keep[t]) continue;
    long long b = cnt[t], e = cnt[(size_t) t + 1];
    GatherList& gl = out[(e - b) > LARGE_FANIN ? 1 : 0];
    gl.src.insert(gl.src.end(), sorted.begin() + b, sorted.begin() + e);
    gl.ptr.push_back((long long) gl.src.size());
    gl.target.push_back(out_index[t]);
  }
}

}  // namespace

void Problem::window(long long mesh_vars, int seg_vars, long long nseg, long long seg_lo, long long seg_hi, int halo) {
  if (analyzed) throw std::logic_error("window: the problem has been analysed already");
  if (seg_vars <= 0 || nseg <= 0 || seg_lo < 0 || seg_hi > nseg || seg_lo >= seg_hi || mesh_vars > nvars)
    throw std::invalid_argument("window: bad segment range");
  win = Window();
  win.on = true;
  win.mesh_vars = mesh_vars;
  win.seg_vars = seg_vars;
  win.nseg = nseg;
  win.seg_lo = seg_lo;
  win.seg_hi = seg_hi;
  win.halo = halo;
  win.first = seg_lo == 0;
  win.last = seg_hi == nseg;
  const long long k_lo = std::max<long long>(seg_lo - halo, 0), k_hi = std::min<long long>(seg_hi + halo, nseg);
  const size_t nf = funcs.size();
  win.app_base.assign(nf, 0);
  win.nappl_whole.assign(nf, 0);
  win.own_lo.assign(nf, 0);
  win.own_hi.assign(nf, 0);
  for (size_t f = 0; f < nf; f++) {
    FuncRec& fn = funcs[f];
    win.nappl_whole[f] = fn.nappl;
    int keep_lo = -1, keep_hi = -1, own_lo = -1, own_hi = -1;
    // inputs that are the same variable in every application do not place an application on the mesh
    std::vector<char> fixed((size_t) fn.ir, 0);
    if (fn.nappl > 1)
      for (int i = 0; i < fn.ir; i++) {
        fixed[(size_t) i] = 1;
        const int v0 = fn.vindex[(size_t) i];
        for (int a = 1; a < fn.nappl && fixed[(size_t) i]; a++)
          if (fn.vindex[(size_t) a * fn.ir + i] != v0) fixed[(size_t) i] = 0;
        if (fixed[(size_t) i] && v0 < mesh_vars) win.far_mesh_vars.push_back(v0);
      }
    for (int a = 0; a < fn.nappl; a++) {
      long long vmin = -1;
      for (int i = 0; i < fn.ir; i++) {
        if (fixed[(size_t) i]) continue;
        const long long v = fn.vindex[(size_t) a * fn.ir + i];
        if (v < mesh_vars && (vmin < 0 || v < vmin)) vmin = v;
      }
      const long long home = vmin < 0 ? -1 : std::min(vmin / seg_vars, nseg - 1);
      const bool keep = home < 0 || (home >= k_lo && home < k_hi);
      const bool own = home < 0 ? win.last : (home >= seg_lo && home < seg_hi);
      if (keep) {
        if (keep_lo < 0) keep_lo = a;
        else if (keep_hi != a) throw std::runtime_error("window: kept applications of " + fn.name + " are not contiguous");
        keep_hi = a + 1;
      }
      if (own) {
        if (own_lo < 0) own_lo = a;
        else if (own_hi != a) throw std::runtime_error("window: owned applications of " + fn.name + " are not contiguous");
        own_hi = a + 1;
      }
    }
    if (keep_lo < 0) keep_lo = keep_hi = 0;
    if (own_lo < 0) own_lo = own_hi = keep_lo;
    win.app_base[f] = keep_lo;
    win.own_lo[f] = own_lo - keep_lo;
    win.own_hi[f] = own_hi - keep_lo;
    const int n = keep_hi - keep_lo;
    fn.vindex = std::vector<int>(fn.vindex.begin() + (size_t) keep_lo * fn.ir, fn.vindex.begin() + (size_t) keep_hi * fn.ir);
    if (fn.role != 0)
      fn.cindex = std::vector<int>(fn.cindex.begin() + (size_t) keep_lo * fn.orows, fn.cindex.begin() + (size_t) keep_hi * fn.orows);
    if (fn.nappconst > 0)
      fn.appconst = std::vector<double>(fn.appconst.begin() + (size_t) keep_lo * fn.nappconst,
                                        fn.appconst.begin() + (size_t) keep_hi * fn.nappconst);
    fn.nappl = n;
  }
  std::sort(win.far_mesh_vars.begin(), win.far_mesh_vars.end());
  win.far_mesh_vars.erase(std::unique(win.far_mesh_vars.begin(), win.far_mesh_vars.end()), win.far_mesh_vars.end());
}

void Problem::analyze() {
  const int PV = nvars, SV = niq, EC = neq, IC = niq;
  kktdim = PV + SV + EC + IC;
  const int eqoffset = PV + SV, iqoffset = PV + SV + EC;

  // evaluation order: objectives, equalities, inequalities (Threads = 1 order of getMATSpace)
  order.clear();
  for (int role = 0; role < 3; role++)
    for (int i = 0; i < (int) funcs.size(); i++)
      if (funcs[i].role == role) order.push_back(i);

  trip_base.assign(funcs.size(), 0);
  fdest_base.assign(funcs.size(), 0);
  n_user = 0;
  long long nf = 0;
  for (int fi: order) {
    FuncRec& fn = funcs[fi];
    if ((long long) fn.vindex.size() != (long long) fn.nappl * fn.ir) throw std::invalid_argument("analyze: bad vindex size in " + fn.name);
    if (fn.role != 0 && (long long) fn.cindex.size() != (long long) fn.nappl * fn.orows)
      throw std::invalid_argument("analyze: bad cindex size in " + fn.name);
    for (int v: fn.vindex)
      if (v < 0 || v >= PV) throw std::out_of_range("analyze: variable index out of range in " + fn.name);
    const int nc = (fn.role == 1) ? EC : IC;
    if (fn.role != 0)
      for (int c: fn.cindex)
        if (c < 0 || c >= nc) throw std::out_of_range("analyze: constraint index out of range in " + fn.name);
    trip_base[fi] = n_user;
    n_user += (long long) fn.nappl * fn.triplets_per_app();
    fdest_base[fi] = nf;
    if (fn.role != 0) nf += (long long) fn.nappl * fn.orows;
  }
  n_solver = (long long) SV + PV + SV + EC + IC;
  const long long T = n_user + n_solver;
  if (T >= (1ll << 31)) throw std::length_error("analyze: more than 2^31 KKT triplets");

  auto conoff = [&](const FuncRec& fn) { return fn.role == 1 ? eqoffset : (fn.role == 2 ? iqoffset : 0); };
  // solver-owned entries in the reference's storage order (NonLinearProgram.cpp:198-265)
  // A windowed problem (Problem::window) keeps the whole problem's dimensions and the numbering of the solver's
  // coefficients, but creates matrix entries only in the rows its applications touch: everything here then costs in
  // proportion to the window, not to the phase (a 1.6 M-segment phase has 1.3e8 solver-owned entries).
  std::vector<unsigned char> keep_v, keep_e, keep_i;
  if (win.on) {
    keep_v.assign((size_t) std::max(PV, 1), 0);
    keep_e.assign((size_t) std::max(EC, 1), 0);
    keep_i.assign((size_t) std::max(IC, 1), 0);
    for (const FuncRec& fn: funcs) {
      for (int v: fn.vindex) keep_v[(size_t) v] = 1;
      if (fn.role != 0)
        for (int c: fn.cindex) (fn.role == 1 ? keep_e : keep_i)[(size_t) c] = 1;
    }
  }
  auto keep_var = [&](int i) { return !win.on || keep_v[(size_t) i]; };
  auto keep_eq = [&](int i) { return !win.on || keep_e[(size_t) i]; };
  auto keep_iq = [&](int i) { return !win.on || keep_i[(size_t) i]; };
  // cb(row, col, kept): entries that are not kept still count in the numbering of the solver's coefficients
  auto for_each_solver = [&](auto&& cb) {
    for (int i = 0; i < SV; i++) cb(PV + SV + EC + i, PV + i, keep_iq(i));  // slack Jacobian ones
    for (int i = 0; i < PV; i++) cb(i, i, keep_var(i));                      // primal diagonal
    for (int i = 0; i < SV; i++) cb(PV + i, PV + i, keep_iq(i));            // slack diagonal
    for (int i = 0; i < EC; i++) cb(PV + SV + i, PV + SV + i, keep_eq(i));  // equality pivots
    for (int i = 0; i < IC; i++) cb(PV + SV + EC + i, PV + SV + EC + i, keep_iq(i));
  };

  // pass 1: count per CSR row (upper triangle: row = min, col = max)
  std::vector<long long> rowstart((size_t) kktdim + 1, 0);
  auto count_cb = [&](int r, int c) { rowstart[(size_t) std::min(r, c) + 1]++; };
  for (int fi: order) for_each_triplet(funcs[fi], conoff(funcs[fi]), count_cb);
  long long n_solver_kept = 0;
  for_each_solver([&](int r, int c, bool kept) {
    if (!kept) return;
    count_cb(r, c);
    n_solver_kept++;
  });
  for (int i = 0; i < kktdim; i++) rowstart[(size_t) i + 1] += rowstart[i];
  // pass 2: bucket the columns
  std::vector<int> colbuf((size_t) (n_user + n_solver_kept));
  {
    std::vector<long long> pos(rowstart.begin(), rowstart.end() - 1);
    auto fill_cb = [&](int r, int c) { colbuf[(size_t) pos[std::min(r, c)]++] = std::max(r, c); };
    for (int fi: order) for_each_triplet(funcs[fi], conoff(funcs[fi]), fill_cb);
    for_each_solver([&](int r, int c, bool kept) {
      if (kept) fill_cb(r, c);
    });
  }
  // pass 3: sort + unique per row (parallel over rows)
  std::vector<int> rowcnt(kktdim, 0);
  {
    unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++) {
      th.emplace_back([&, t]() {
        int lo = (int) ((long long) kktdim * t / nt), hi = (int) ((long long) kktdim * (t + 1) / nt);
        for (int r = lo; r < hi; r++) {
          int* b = colbuf.data() + rowstart[r];
          int* e = colbuf.data() + rowstart[(size_t) r + 1];
          std::sort(b, e);
          rowcnt[r] = (int) (std::unique(b, e) - b);
        }
      });
    }
    for (auto& x: th) x.join();
  }
  long long nnz_ll = 0;
  for (int r = 0; r < kktdim; r++) nnz_ll += rowcnt[r];
  if (nnz_ll >= (1ll << 31)) throw std::length_error("analyze: CSR nnz exceeds int32");
  indptr.assign((size_t) kktdim + 1, 0);
  indices.resize((size_t) nnz_ll);
  for (int r = 0; r < kktdim; r++) {
    indptr[(size_t) r + 1] = indptr[r] + rowcnt[r];
    std::copy(colbuf.begin() + rowstart[r], colbuf.begin() + rowstart[r] + rowcnt[r], indices.begin() + indptr[r]);
  }
  std::vector<int>().swap(colbuf);

  // pass 4: KKTLocations (value-array offset of every triplet)
  kktloc.resize((size_t) T);
  {
    long long k = 0;
    auto loc_cb = [&](int r, int c) {
      int rr = std::min(r, c), cc = std::max(r, c);
      const int* b = indices.data() + indptr[rr];
      const int* e = indices.data() + indptr[(size_t) rr + 1];
      kktloc[(size_t) k++] = (int) (std::lower_bound(b, e, cc) - indices.data());
    };
    for (int fi: order) for_each_triplet(funcs[fi], conoff(funcs[fi]), loc_cb);
    for_each_solver([&](int r, int c, bool kept) {
      if (kept) loc_cb(r, c);
      else kktloc[(size_t) k++] = -1;  // no entry in this window's pattern
    });
  }

  // pass 5: classify slots.  A slot written by exactly one user triplet and no solver entry is stored
  // directly by the evaluating lane; everything else goes through the fixed-order gather.
  const int nnz_i = (int) nnz_ll;
  std::vector<unsigned char> nsrc((size_t) nnz_i, 0);  // saturating count
  for (long long k = 0; k < T; k++) {
    if (kktloc[(size_t) k] < 0) continue;
    unsigned char& c = nsrc[(size_t) kktloc[(size_t) k]];
    if (k >= n_user)
      c = 2;
    else if (c < 2)
      c++;
  }
  kkt_dest.resize((size_t) n_user);
  long long stage_pos = 0;
  std::vector<int> tgt_of_src, stage_of_src;
  for (long long k = 0; k < n_user; k++) {
    int s = kktloc[(size_t) k];
    if (nsrc[(size_t) s] == 1) {
      kkt_dest[(size_t) k] = s;
    } else {
      kkt_dest[(size_t) k] = ~(int) stage_pos;
      tgt_of_src.push_back(s);
      stage_of_src.push_back((int) stage_pos);
      stage_pos++;
    }
  }
  stage_solver = stage_pos;
  for (long long k = 0; k < n_solver; k++) {
    if (kktloc[(size_t) (n_user + k)] >= 0) {
      tgt_of_src.push_back(kktloc[(size_t) (n_user + k)]);
      stage_of_src.push_back((int) stage_pos);
    }
    stage_pos++;  // the coefficient keeps its slot whether or not this window has an entry for it
  }
  {
    std::vector<char> keep((size_t) nnz_i);
    std::vector<int> outidx((size_t) nnz_i);
    for (int s = 0; s < nnz_i; s++) {
      keep[s] = nsrc[s] >= 2;
      outidx[s] = s;
    }
    build_gather(tgt_of_src, stage_of_src, nnz_i, outidx, keep, g_kkt);
  }

  // gradients: every application writes its IR gradient entries to private stage slots (the
  // reference's RHScoeffs, IndexingData.h:96-105), summed per variable in source order.
  g_stage.assign(funcs.size(), -1);
  v_stage.assign(funcs.size(), -1);
  auto grad_gather = [&](int role_lo, int role_hi, GatherList out[2]) {
    std::vector<int> tg, st;
    for (int fi: order) {
      FuncRec& fn = funcs[fi];
      if (fn.role < role_lo || fn.role > role_hi) continue;
      g_stage[fi] = stage_pos;
      for (long long q = 0; q < (long long) fn.nappl * fn.ir; q++) {
        tg.push_back(fn.vindex[(size_t) q]);
        st.push_back((int) (stage_pos + q));
      }
      stage_pos += (long long) fn.nappl * fn.ir;
    }
    std::vector<char> keep((size_t) PV, 1);
    std::vector<int> outidx((size_t) PV);
    for (int i = 0; i < PV; i++) outidx[i] = i;
    build_gather(tg, st, PV, outidx, keep, out);
  };
  grad_gather(0, 0, g_pgx);
  grad_gather(1, 2, g_agx);

  // objective value
  {
    std::vector<int> tg, st;
    for (int fi: order) {
      FuncRec& fn = funcs[fi];
      if (fn.role != 0) continue;
      v_stage[fi] = stage_pos;
      for (int a = 0; a < fn.nappl; a++) {
        tg.push_back(0);
        st.push_back((int) (stage_pos + a));
      }
      stage_pos += fn.nappl;
    }
    std::vector<char> keep(1, 1);
    std::vector<int> outidx(1, 0);
    build_gather(tg, st, 1, outidx, keep, g_val);
  }

  // constraint values: unique rows are stored directly; accumulated rows are staged and summed
  f_dest.assign((size_t) nf, 0);
  for (int role = 1; role <= 2; role++) {
    const int nc = (role == 1) ? EC : IC;
    std::vector<int> users((size_t) nc, 0);
    for (int fi: order)
      if (funcs[fi].role == role)
        for (int c: funcs[fi].cindex) users[(size_t) c]++;
    std::vector<int> tg, st;
    for (int fi: order) {
      FuncRec& fn = funcs[fi];
      if (fn.role != role) continue;
      for (size_t q = 0; q < fn.cindex.size(); q++) {
        int c = fn.cindex[q];
        if (users[(size_t) c] == 1) {
          f_dest[(size_t) fdest_base[fi] + q] = c;
        } else {
          f_dest[(size_t) fdest_base[fi] + q] = ~(int) stage_pos;
          tg.push_back(c);
          st.push_back((int) stage_pos);
          stage_pos++;
        }
      }
    }
    std::vector<char> keep((size_t) nc);
    std::vector<int> outidx((size_t) nc);
    for (int c = 0; c < nc; c++) {
      keep[c] = users[c] != 1;  // rows nobody writes are zero-filled by the gather
      outidx[c] = c;
    }
    build_gather(tg, st, nc, outidx, keep, role == 1 ? g_fxe : g_fxi);
  }
  if (stage_pos >= (1ll << 31)) throw std::length_error("analyze: staging buffer exceeds int32");
  stage_size = stage_pos;
  analyzed = true;
}

// ---------------------------------------------------------------------------------------------------------------
// Multi-GPU partition (see RankPartition in nlp.hpp)
// ---------------------------------------------------------------------------------------------------------------
RankPartition make_partition(const Problem& prob, int rank, int world) {
  if (!prob.analyzed) throw std::invalid_argument("make_partition: analyze() first");
  if (world < 1 || world > 64 || rank < 0 || rank >= world) throw std::invalid_argument("make_partition: bad rank / world");
  RankPartition P;
  P.rank = rank;
  P.world = world;
  const size_t nf = prob.funcs.size();
  P.app_begin.resize(nf);
  P.app_end.resize(nf);
  auto cut = [&](int nappl, int r) { return (int) ((long long) nappl * r / world); };
  for (size_t f = 0; f < nf; f++) {
    P.app_begin[f] = cut(prob.funcs[f].nappl, rank);
    P.app_end[f] = cut(prob.funcs[f].nappl, rank + 1);
  }
  // rank evaluating application `a` of a function with `n` applications: the r with cut(n, r) <= a < cut(n, r + 1)
  auto rank_of = [&](int n, int a) {
    int r = (int) (((long long) a * world + world - 1) / std::max(n, 1));
    r = std::min(std::max(r, 0), world - 1);
    while (r > 0 && cut(n, r) > a) r--;
    while (r + 1 < world && cut(n, r + 1) <= a) r++;
    return (unsigned char) r;
  };
  // which rank writes each staging slot (255: solver-owned coefficient, follows the owner of its target)
  std::vector<unsigned char> slot_rank((size_t) prob.stage_size, 255);
  for (int fi: prob.order) {
    const FuncRec& fn = prob.funcs[fi];
    const int E = fn.triplets_per_app();
    const long long tb = prob.trip_base[fi];
    for (int a = 0; a < fn.nappl; a++) {
      const unsigned char r = rank_of(fn.nappl, a);
      for (int e = 0; e < E; e++) {
        const int d = prob.kkt_dest[(size_t) (tb + (long long) a * E + e)];
        if (d < 0) slot_rank[(size_t) ~d] = r;
      }
      if (prob.g_stage[fi] >= 0)
        for (int i = 0; i < fn.ir; i++) slot_rank[(size_t) (prob.g_stage[fi] + (long long) a * fn.ir + i)] = r;
      if (prob.v_stage[fi] >= 0) slot_rank[(size_t) (prob.v_stage[fi] + a)] = r;
      if (fn.role != 0)
        for (int o = 0; o < fn.orows; o++) {
          const int d = prob.f_dest[(size_t) (prob.fdest_base[fi] + (long long) a * fn.orows + o)];
          if (d < 0) slot_rank[(size_t) ~d] = r;
        }
    }
  }
  const GatherList* src_lists[OA_COUNT] = {prob.g_kkt, prob.g_pgx, prob.g_agx, prob.g_fxe, prob.g_fxi, prob.g_val};
  for (int a = 0; a < OA_COUNT; a++) {
    for (int w = 0; w < 2; w++) {
      P.g[a][w].ptr.assign(1, 0);
      P.g[a][w].src.clear();
      P.g[a][w].target.clear();
    }
    for (int w = 0; w < 2; w++) {
      const GatherList& gl = src_lists[a][w];
      const int nt = gl.ntargets();
      for (int t = 0; t < nt; t++) {
        const long long b = gl.ptr[t], e = gl.ptr[(size_t) t + 1];
        unsigned long long mask = 0;
        for (long long k = b; k < e; k++) {
          const unsigned char r = slot_rank[(size_t) gl.src[(size_t) k]];
          if (r != 255) mask |= 1ull << r;
        }
        int owner;
        if (mask == 0)
          owner = (int) ((long long) t * world / std::max(nt, 1));  // solver-only or empty target: spread by position
        else
          owner = __builtin_ctzll(mask);
        const bool multi = (mask & (mask - 1)) != 0;
        const bool mine = multi ? ((mask >> rank) & 1ull) != 0 : owner == rank;
        if (multi) {
          P.cut_target[a].push_back(gl.target[t]);
          P.cut_owner[a].push_back((unsigned char) owner);
          P.cut_mine[a].push_back(mine ? 1 : 0);
        }
        if (!mine) continue;
        // own sources (and the solver coefficient if this rank owns the target), in the original fixed order; the list
        // keeps its fan-in class so the same kernels serve it
        GatherList& out = P.g[a][w];
        for (long long k = b; k < e; k++) {
          const int sidx = gl.src[(size_t) k];
          const unsigned char r = slot_rank[(size_t) sidx];
          if (r == rank || (r == 255 && owner == rank)) out.src.push_back(sidx);
        }
        out.ptr.push_back((long long) out.src.size());
        out.target.push_back(gl.target[t]);
      }
    }
  }
  return P;
}

std::vector<std::pair<long long, long long>> Problem::input_ranges(int which, long long gap) const {
  if (which < 0 || which > 2) throw std::invalid_argument("input_ranges: which must be 0 (X), 1 (LE) or 2 (LI)");
  const long long n = which == 0 ? nvars : which == 1 ? neq : niq;
  std::vector<std::pair<long long, long long>> out;
  if (n <= 0) return out;
  if (!win.on) {
    out.emplace_back(0, n);
    return out;
  }
  std::vector<unsigned char> used((size_t) n, 0);
  for (const FuncRec& fr: funcs) {
    if (which == 0) {
      for (int v: fr.vindex) used.at((size_t) v) = 1;
    } else if (fr.role == which) {
      for (int c: fr.cindex) used.at((size_t) c) = 1;
    }
  }
  long long i = 0;
  while (i < n) {
    if (!used[(size_t) i]) {
      i++;
      continue;
    }
    long long hi = i + 1, last = i;  // grow while the next used element is within `gap`
    while (hi < n && hi - last <= gap) {
      if (used[(size_t) hi]) last = hi;
      hi++;
    }
    out.emplace_back(i, last + 1);
    i = last + 1;
  }
  return out;
}

}  // namespace ab
