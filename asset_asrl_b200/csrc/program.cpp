// Scalar DAG -> device program: FMA fusion, depth-first ordering (short live ranges), bundling of independent
// operations and lane-private slot allocation.  See the Program comment in expr.hpp for the execution model.
#include <algorithm>
#include <cstring>
#include <map>
#include <stdexcept>

#include "expr.hpp"

namespace ab {

int triplet_offset_h(int i, int j, int ir, int orows, bool lay_h, bool lay_j) {
  // column i starts after the i previous columns: each has (ir - c) Hessian rows and orows Jacobian rows
  int base = 0;
  if (lay_h) base += i * ir - (i * (i - 1)) / 2;
  if (lay_j) base += i * orows;
  return base + (j - i);
}
int triplet_offset_j(int i, int k, int ir, int orows, bool lay_h, bool lay_j) {
  int base = 0;
  if (lay_h) base += i * ir - (i * (i - 1)) / 2 + (ir - i);
  if (lay_j) base += i * orows;
  return base + k;
}

namespace {

// An operand of a machine operation: a DAG node (value produced by another machine op or a constant) + negation.
struct Opnd {
  int node = -1;  // DAG node id; constants are resolved through the node's imm
  bool neg = false;
};
struct MOp {
  uint32_t kind = PK_FMA;
  int node = -1;          // DAG node this op defines (-1 for OUT)
  Opnd a, b, c;           // FMA: (+-a)*b + (+-c);  SPECIAL: op(a, b[, c]);  OUT: value a
  uint16_t sop = 0;       // SPECIAL: ab::Op
  double imm = 0.0;       // SPECIAL POWC exponent
  uint32_t x0 = 0, x1 = 0;  // LOAD: index, type;  OUT: out kind, index
};

struct Builder {
  ScalarFn& f;
  SDag& dag;
  std::vector<int> uses;      // use count of each node by needed nodes / outputs
  std::vector<char> emitted;
  std::vector<MOp> ops;
  int n_one, n_zero;

  explicit Builder(ScalarFn& f_) : f(f_), dag(f_.dag) {
    n_one = dag.constant(1.0);
    n_zero = dag.constant(0.0);
  }
  bool is_product(int n) const {
    const uint16_t op = dag.n[n].op;
    return op == OP_MUL || op == OP_MULC || op == OP_SQUARE || op == OP_NEG;
  }
  // a product node as (+-p) * q
  void product_parts(int n, Opnd& p, Opnd& q) {
    const SNode& nd = dag.n[n];
    switch (nd.op) {
      case OP_MUL: p = {nd.a, false}; q = {nd.b, false}; break;
      case OP_MULC: p = {nd.a, false}; q = {dag.constant(nd.imm), false}; break;
      case OP_SQUARE: p = {nd.a, false}; q = {nd.a, false}; break;
      default: p = {nd.a, true}; q = {n_one, false}; break;  // NEG
    }
  }
  void count_uses(const std::vector<int>& roots) {
    uses.assign(dag.n.size(), 0);
    std::vector<char> seen(dag.n.size(), 0);
    std::vector<int> st;
    for (int r: roots) {
      uses[r]++;
      st.push_back(r);
    }
    while (!st.empty()) {
      int i = st.back();
      st.pop_back();
      if (seen[i]) continue;
      seen[i] = 1;
      const SNode& nd = dag.n[i];
      if (nd.op == OP_CONST || nd.op == OP_INPUT || nd.op == OP_LAMBDA || nd.op == OP_APPCONST) continue;
      for (int o: {nd.a, nd.b, nd.c})
        if (o >= 0) {
          uses[o]++;
          st.push_back(o);
        }
    }
  }
  bool fusable(int n) const { return is_product(n) && uses[n] == 1 && !dag.is_const(n); }

  // emit everything node n needs, then n itself (post-order)
  void visit(int n) {
    if (dag.is_const(n) || emitted[n]) return;
    // iterative DFS would be safer for very deep DAGs; depth here is bounded by expression depth (hundreds)
    const SNode nd = dag.n[n];
    MOp m;
    m.node = n;
    auto dep = [&](const Opnd& o) {
      if (o.node >= 0) visit(o.node);
    };
    switch (nd.op) {
      case OP_INPUT: m.kind = PK_LOAD; m.x0 = (uint32_t) nd.a; m.x1 = PL_X; break;
      case OP_LAMBDA: m.kind = PK_LOAD; m.x0 = (uint32_t) nd.a; m.x1 = PL_LAMBDA; break;
      case OP_APPCONST: m.kind = PK_LOAD; m.x0 = (uint32_t) nd.a; m.x1 = PL_APPCONST; break;
      case OP_ADD:
      case OP_SUB: {
        const bool sub = nd.op == OP_SUB;
        const int x = nd.a, y = nd.b;  // x + y or x - y
        if (fusable(y)) {
          product_parts(y, m.a, m.b);
          m.a.neg ^= sub;
          m.c = {x, false};
        } else if (fusable(x)) {
          product_parts(x, m.a, m.b);
          m.c = {y, sub};
        } else {
          m.a = {x, false};
          m.b = {n_one, false};
          m.c = {y, sub};
        }
      } break;
      case OP_ADDC: {
        const int k = dag.constant(nd.imm);
        if (fusable(nd.a)) {
          product_parts(nd.a, m.a, m.b);
        } else {
          m.a = {nd.a, false};
          m.b = {n_one, false};
        }
        m.c = {k, false};
      } break;
      case OP_MUL:
      case OP_MULC:
      case OP_SQUARE:
      case OP_NEG:
        product_parts(n, m.a, m.b);
        m.c = {n_zero, false};
        break;
      case OP_FMA:
        m.a = {nd.a, false};
        m.b = {nd.b, false};
        m.c = {nd.c, false};
        break;
      default:
        m.kind = PK_SPECIAL;
        m.sop = nd.op;
        m.imm = nd.imm;
        m.a = {nd.a, false};
        if (nd.b >= 0) m.b = {nd.b, false};
        if (nd.c >= 0) m.c = {nd.c, false};
        break;
    }
    if (m.kind != PK_LOAD) {
      dep(m.a);
      dep(m.b);
      dep(m.c);
    }
    emitted[n] = 1;
    ops.push_back(m);
  }
  void out(int node, uint32_t okind, uint32_t index) {
    visit(node);
    MOp m;
    m.kind = PK_OUT;
    m.a = {node, false};
    m.x0 = okind;
    m.x1 = index;
    ops.push_back(m);
  }
};

}  // namespace

Program emit_program(ScalarFn& f, const ProgramSpec& spec) {
  if ((spec.want_g || spec.want_h) && (int) f.grads.size() != f.nin) f.make_gradient();
  if ((spec.want_j || spec.want_h) && f.jac.empty()) f.make_second_order();
  const int ir = f.nin, orr = f.nout;
  const bool lay_h = !spec.linear, lay_j = !spec.is_objective;

  // ---- outputs in the triplet order of DenseFunctionBase::getKKTSpace: F, then per column i: H rows j >= i, J rows, G_i
  std::vector<ProgOut> outs;
  if (spec.want_f)
    for (int k = 0; k < orr; k++) outs.push_back({f.outs[k], PO_F, (uint32_t) k});
  for (int i = 0; i < ir; i++) {
    if (spec.want_h && lay_h)
      for (int j = i; j < ir; j++)
        outs.push_back({f.hess[(size_t) j * ir + i], PO_H, (uint32_t) triplet_offset_h(i, j, ir, orr, lay_h, lay_j)});
    if (spec.want_j && lay_j)
      for (int k = 0; k < orr; k++)
        outs.push_back({f.jac[(size_t) k * ir + i], PO_J, (uint32_t) triplet_offset_j(i, k, ir, orr, lay_h, lay_j)});
    if (spec.want_g) outs.push_back({f.grads[i], PO_G, (uint32_t) i});
  }
  return emit_program_outs(f, outs);
}

Program emit_program_outs(ScalarFn& f, const std::vector<ProgOut>& outs) {
  Builder B(f);
  SDag& dag = f.dag;
  const int ir = f.nin, orr = f.nout;
  {
    std::vector<int> roots;
    for (auto& o: outs) roots.push_back(o.node);
    B.count_uses(roots);
    // an output that is itself a fusable product must still be materialised
    for (auto& o: outs) B.uses[o.node] += 1;
  }
  B.emitted.assign(dag.n.size(), 0);
  for (auto& o: outs) B.out(o.node, o.kind, o.index);
  B.emitted.resize(dag.n.size(), 0);
  std::vector<MOp>& ops = B.ops;
  const int n = (int) ops.size();

  // ---- dependency info: producer op of each node
  std::map<int, int> prod;  // node -> op index
  for (int k = 0; k < n; k++)
    if (ops[k].node >= 0) prod[ops[k].node] = k;
  auto opnds = [&](const MOp& m, int out[3]) {
    int c = 0;
    if (m.kind == PK_LOAD) return 0;
    for (const Opnd* o: {&m.a, &m.b, &m.c})
      if (o->node >= 0 && !dag.is_const(o->node)) out[c++] = o->node;
    return c;
  };

  // ---- windowed list scheduling into bundles of <= 4 independent ops of one kind
  constexpr int WINDOW = 48;
  std::vector<int> bundle_of(n, -1);  // closed bundle index per op
  std::vector<std::vector<int>> bundles;
  std::vector<char> done(n, 0);
  int head = 0, nbund = 0;
  while (head < n) {
    if (done[head]) {
      head++;
      continue;
    }
    const uint32_t kind = ops[head].kind;
    const int cap = (kind == PK_SPECIAL) ? 1 : 4;
    std::vector<int> cur;
    for (int k = head; k < n && k < head + WINDOW && (int) cur.size() < cap; k++) {
      if (done[k] || ops[k].kind != kind) continue;
      int on[3];
      const int c = opnds(ops[k], on);
      bool ready = true;
      for (int q = 0; q < c; q++) {
        const int p = prod.at(on[q]);
        if (!done[p] || bundle_of[p] == nbund) ready = false;
      }
      if (!ready) {
        // outputs leave in the order they were asked for (the specialiser stores runs of consecutive outputs)
        if (kind == PK_OUT) break;
        continue;
      }
      cur.push_back(k);
    }
    if (cur.empty()) throw std::logic_error("emit_program: scheduler stalled");
    for (int k: cur) {
      done[k] = 1;
      bundle_of[k] = nbund;
    }
    bundles.push_back(cur);
    nbund++;
  }

  // ---- liveness: last bundle reading each node
  std::map<int, int> last;
  for (int b = 0; b < nbund; b++)
    for (int k: bundles[b]) {
      int on[3];
      const int c = opnds(ops[k], on);
      for (int q = 0; q < c; q++) last[on[q]] = b;
    }
  // ---- constants
  Program P;
  P.nin = ir;
  P.nout = orr;
  std::map<uint64_t, int> cidx;
  auto const_index = [&](double v) {
    uint64_t u;
    std::memcpy(&u, &v, 8);
    auto it = cidx.find(u);
    if (it != cidx.end()) return it->second;
    P.consts.push_back(v);
    cidx.emplace(u, (int) P.consts.size() - 1);
    return (int) P.consts.size() - 1;
  };
  const_index(0.0);  // constant 0 always exists: absent operands and padding instructions read it
  // ---- slot allocation (reads of a bundle happen before its writes, so a slot freed in a bundle can be reused by it)
  std::map<int, int> slot;
  std::vector<int> freelist;
  int nslots = 0;
  int maxapp = -1;
  auto enc = [&](const Opnd& o) -> uint32_t {
    if (o.node < 0) return (uint32_t) const_index(0.0) << 3;
    if (dag.is_const(o.node)) return ((uint32_t) const_index(dag.cval(o.node)) << 3) | (o.neg ? P_NEG : 0u);
    return ((uint32_t) slot.at(o.node) << 3) | P_SLOT | (o.neg ? P_NEG : 0u);
  };
  for (int b = 0; b < nbund; b++) {
    const std::vector<int>& cur = bundles[b];
    PInstr ins[4];
    std::memset(ins, 0, sizeof(ins));
    const uint32_t kind = ops[cur[0]].kind;
    // operands first
    for (size_t q = 0; q < cur.size(); q++) {
      const MOp& m = ops[cur[q]];
      switch (kind) {
        case PK_FMA:
          ins[q].a = enc(m.a);
          ins[q].b = enc(m.b);
          ins[q].c = enc(m.c);
          P.n_fma++;
          break;
        case PK_SPECIAL:
          ins[0].a = enc(m.a);
          ins[0].b = (m.b.node >= 0) ? enc(m.b) : 0u;
          ins[0].c = (uint32_t) m.sop | ((uint32_t) const_index(m.imm) << 8);
          ins[1].a = (m.c.node >= 0) ? enc(m.c) : 0u;
          P.n_special++;
          break;
        case PK_LOAD:
          ins[q].a = m.x0;
          ins[q].b = m.x1;
          if (m.x1 == PL_APPCONST) maxapp = std::max(maxapp, (int) m.x0);
          P.n_load++;
          break;
        case PK_OUT:
          ins[q].a = enc(m.a);
          ins[q].b = m.x0;
          ins[q].c = m.x1;
          P.n_out++;
          break;
        default: break;
      }
    }
    // release operands dying here
    for (int k: cur) {
      int on[3];
      const int c = opnds(ops[k], on);
      for (int q = 0; q < c; q++) {
        auto it = slot.find(on[q]);
        if (it != slot.end() && last.at(on[q]) == b) {
          freelist.push_back(it->second);
          slot.erase(it);
        }
      }
    }
    // destinations
    for (size_t q = 0; q < cur.size(); q++) {
      const MOp& m = ops[cur[q]];
      if (m.node < 0) continue;
      int s;
      if (!freelist.empty()) {
        s = freelist.back();
        freelist.pop_back();
      } else {
        s = nslots++;
      }
      if (last.find(m.node) == last.end()) {
        freelist.push_back(s);  // never read (dead value): slot is free again right away
      } else {
        slot[m.node] = s;
      }
      ins[q].d = ((uint32_t) s << 3) | P_SLOT;
    }
    ins[0].d |= (kind << 28) | ((uint32_t) cur.size() << 24);
    for (int q = 0; q < 4; q++) P.code.push_back(ins[q]);
    P.n_ops += (int) cur.size();
  }
  PInstr endb[4];
  std::memset(endb, 0, sizeof(endb));
  endb[0].d = (uint32_t) PK_END << 28;
  for (int q = 0; q < 4; q++) P.code.push_back(endb[q]);
  P.nslots = std::max(nslots, 1);
  P.nappconst = maxapp + 1;
  if (nslots >= (1 << 21)) throw std::length_error("emit_program: too many slots");
  if (P.n_special) {  // interpolation-table data travel with the program that reads them
    for (const MOp& m: B.ops)
      if (m.kind == PK_SPECIAL && (m.sop == OP_TABIDX || m.sop == OP_TABLOAD)) {
        P.tables = dag.tabpool;
        break;
      }
  }
  return P;
}

// Rewrites operand words from (index << 3 | flags) to (shared-memory byte offset | flags) for a block of `nthreads`
// lane-private slots laid out [slot][thread] after the shared constants, and pads short FMA bundles with
// scratch = 0 * 0 + 0 so the interpreter can always run four operations.
LinkedProgram link_program(const Program& P, unsigned smem_budget, const AffineMaps* am) {
  LinkedProgram L;
  const unsigned const_bytes = (unsigned) ((P.consts.size() * 8 + 15) / 16 * 16);
  const unsigned per_thread = (unsigned) (P.nslots + 1) * 8u;  // +1: scratch slot
  const unsigned code_bytes = 2u * PROG_CHUNK_BUNDLES * 64u;
  if (const_bytes + code_bytes + per_thread * 32u > smem_budget)
    throw std::length_error("function needs " + std::to_string(P.nslots) + " live values per application: too large for the "
                            "shared-memory interpreter (no CPU fallback on this path)");
  unsigned nt = (smem_budget - const_bytes - code_bytes) / per_thread / 32u * 32u;
  nt = std::min(nt, 256u);
  L.nthreads = (int) nt;
  L.smem_bytes = const_bytes + code_bytes + per_thread * nt;
  L.nslots = P.nslots + 1;
  auto lk = [&](uint32_t w) -> uint32_t {
    const uint32_t flags = w & 7u, idx = w >> 3;
    if (flags & P_SLOT) return (const_bytes + code_bytes + idx * nt * 8u) | flags;
    return (idx * 8u) | flags;
  };
  const uint32_t scratch = lk(((uint32_t) P.nslots << 3) | P_SLOT);
  L.code = P.code;
  for (size_t b = 0; b + 3 < L.code.size(); b += 4) {
    PInstr* ins = &L.code[b];
    const uint32_t head = ins[0].d & 0xFF000000u;
    const uint32_t kind = head >> 28, cnt = (head >> 24) & 15u;
    ins[0].d &= 0x00FFFFFFu;
    switch (kind) {
      case PK_FMA:
        for (uint32_t q = 0; q < 4; q++) {
          if (q < cnt) {
            ins[q].d = lk(ins[q].d);
            ins[q].a = lk(ins[q].a);
            ins[q].b = lk(ins[q].b);
            ins[q].c = lk(ins[q].c);
          } else {
            ins[q].d = scratch;
            ins[q].a = ins[q].b = ins[q].c = 0u;
          }
        }
        break;
      case PK_LOAD:
        // linked form: a = index | type << 28, b = affine base, c = affine stride (per application, from app a0)
        for (uint32_t q = 0; q < cnt; q++) {
          ins[q].d = lk(ins[q].d);
          const uint32_t idx = ins[q].a, ty = ins[q].b;
          ins[q].a = idx | (ty << 28);
          ins[q].b = ins[q].c = 0u;
          if (am && am->a1 > am->a0) {
            if (ty == PL_X) {
              ins[q].b = (uint32_t) am->vbase.at(idx);
              ins[q].c = (uint32_t) am->vstride.at(idx);
            } else if (ty == PL_LAMBDA && !am->cbase.empty()) {
              ins[q].b = (uint32_t) am->cbase.at(idx);
              ins[q].c = (uint32_t) am->cstride.at(idx);
            }
          }
        }
        break;
      case PK_OUT:
        // linked form: b = out kind | index << 4, c = affine destination base, d[23:0] = signed affine stride
        for (uint32_t q = 0; q < cnt; q++) {
          ins[q].a = lk(ins[q].a);
          const uint32_t ok = ins[q].b, idx = ins[q].c;
          ins[q].b = ok | (idx << 4);
          int base = 0, stride = 0;
          if (am && am->a1 > am->a0) {
            if (ok == PO_F && !am->fbase.empty()) {
              base = am->fbase.at(idx);
              stride = am->fstride.at(idx);
            } else if (ok == PO_J || ok == PO_H) {
              base = am->kbase.at(idx);
              stride = am->kstride.at(idx);
            }
          }
          ins[q].c = (uint32_t) base;
          ins[q].d = (ins[q].d & 0xFF000000u) | ((uint32_t) stride & 0x00FFFFFFu);
        }
        break;
      case PK_SPECIAL:
        ins[0].d = lk(ins[0].d);
        ins[0].a = lk(ins[0].a);
        ins[0].b = lk(ins[0].b);
        ins[1].a = lk(ins[1].a);
        break;
      default: break;
    }
    if (kind != PK_OUT && (ins[0].d & 0xFF000000u)) throw std::length_error("link_program: shared-memory offset exceeds 24 bits");
    ins[0].d = (ins[0].d & 0x00FFFFFFu) | head;
  }
  return L;
}

}  // namespace ab
