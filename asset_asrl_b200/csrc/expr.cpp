#include "expr.hpp"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <sstream>
#include <stdexcept>

namespace ab {

// ---------------------------------------------------------------------------------------------
// SDag
// ---------------------------------------------------------------------------------------------
static inline uint64_t dbits(double v) {
  uint64_t u;
  std::memcpy(&u, &v, 8);
  return u;
}

int SDag::raw(uint16_t op, int a, int b, int c, double imm) {
  if (imm == 0.0) imm = 0.0;  // fold -0.0 into +0.0 for consing
  Key k {op, a, b, c, dbits(imm)};
  auto it = cons.find(k);
  if (it != cons.end()) return it->second;
  n.push_back(SNode {op, a, b, c, imm});
  int id = (int) n.size() - 1;
  cons.emplace(k, id);
  return id;
}

int SDag::add(int a, int b) {
  if (is_const(a) && is_const(b)) return constant(cval(a) + cval(b));
  if (is_const(a)) std::swap(a, b);
  if (is_const(b)) return addc(a, cval(b));
  if (a > b) std::swap(a, b);
  return raw(OP_ADD, a, b, -1, 0.0);
}
int SDag::sub(int a, int b) {
  if (is_const(a) && is_const(b)) return constant(cval(a) - cval(b));
  if (is_const(b)) return addc(a, -cval(b));
  if (is_const(a)) return addc(neg(b), cval(a));
  if (a == b) return constant(0.0);
  return raw(OP_SUB, a, b, -1, 0.0);
}
int SDag::mul(int a, int b) {
  if (is_const(a) && is_const(b)) return constant(cval(a) * cval(b));
  if (is_const(a)) std::swap(a, b);
  if (is_const(b)) return mulc(a, cval(b));
  if (a == b) return raw(OP_SQUARE, a, -1, -1, 0.0);
  if (a > b) std::swap(a, b);
  return raw(OP_MUL, a, b, -1, 0.0);
}
int SDag::div(int a, int b) {
  if (is_const(a) && is_const(b)) return constant(cval(a) / cval(b));
  if (is_const(b)) return mulc(a, 1.0 / cval(b));
  if (is_const(a)) {
    if (cval(a) == 0.0) return constant(0.0);
    return mulc(unary(OP_INV, b), cval(a));
  }
  return raw(OP_DIV, a, b, -1, 0.0);
}
int SDag::neg(int a) {
  if (is_const(a)) return constant(-cval(a));
  if (n[a].op == OP_NEG) return n[a].a;
  if (n[a].op == OP_MULC) return mulc(n[a].a, -n[a].imm);
  return raw(OP_NEG, a, -1, -1, 0.0);
}
int SDag::addc(int a, double c) {
  if (is_const(a)) return constant(cval(a) + c);
  if (c == 0.0) return a;
  if (n[a].op == OP_ADDC) return addc(n[a].a, n[a].imm + c);
  return raw(OP_ADDC, a, -1, -1, c);
}
int SDag::mulc(int a, double c) {
  if (is_const(a)) return constant(cval(a) * c);
  if (c == 1.0) return a;
  if (c == 0.0) return constant(0.0);
  if (c == -1.0) return neg(a);
  if (n[a].op == OP_MULC) return mulc(n[a].a, n[a].imm * c);
  if (n[a].op == OP_NEG) return mulc(n[a].a, -c);
  return raw(OP_MULC, a, -1, -1, c);
}

static double eval_unary(uint16_t op, double x, double imm) {
  switch (op) {
    case OP_SQUARE: return x * x;
    case OP_INV: return 1.0 / x;
    case OP_SQRT: return std::sqrt(x);
    case OP_EXP: return std::exp(x);
    case OP_LOG: return std::log(x);
    case OP_SIN: return std::sin(x);
    case OP_COS: return std::cos(x);
    case OP_TAN: return std::tan(x);
    case OP_ASIN: return std::asin(x);
    case OP_ACOS: return std::acos(x);
    case OP_ATAN: return std::atan(x);
    case OP_SINH: return std::sinh(x);
    case OP_COSH: return std::cosh(x);
    case OP_TANH: return std::tanh(x);
    case OP_ASINH: return std::asinh(x);
    case OP_ACOSH: return std::acosh(x);
    case OP_ATANH: return std::atanh(x);
    case OP_ABS: return std::fabs(x);
    case OP_SIGN: return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : 0.0);
    case OP_POWC: return std::pow(x, imm);
    default: throw std::logic_error("eval_unary: bad op");
  }
}

int SDag::unary(uint16_t op, int a, double imm) {
  if (is_const(a)) return constant(eval_unary(op, cval(a), imm));
  if (op == OP_POWC) {
    if (imm == 1.0) return a;
    if (imm == 0.0) return constant(1.0);
    if (imm == 2.0) return raw(OP_SQUARE, a, -1, -1, 0.0);
    if (imm == -1.0) return raw(OP_INV, a, -1, -1, 0.0);
    if (imm == 0.5) return raw(OP_SQRT, a, -1, -1, 0.0);
  }
  if (op == OP_NEG) return neg(a);
  return raw(op, a, -1, -1, imm);
}
// ---------------------------------------------------------------------------------------------
// interpolation-table pool
// ---------------------------------------------------------------------------------------------
size_t SDag::tab_add(const double* p, size_t len) {
  for (const TabEntry& e: tabs)
    if (e.len == len && std::memcmp(&tabpool[e.off], p, len * sizeof(double)) == 0) return e.off;
  const size_t off = tabpool.size();
  tabpool.insert(tabpool.end(), p, p + len);
  tabs.push_back({off, len});
  return off;
}
const SDag::TabEntry& SDag::tab_find(size_t off) const {
  for (const TabEntry& e: tabs)
    if (off >= e.off && off < e.off + e.len) return e;
  throw std::logic_error("table offset outside every table");
}
double SDag::tab_import(const SDag& src, double imm) {
  const size_t off = (size_t) imm;
  const TabEntry& e = src.tab_find(off);
  std::vector<double> copy(src.tabpool.begin() + (long) e.off, src.tabpool.begin() + (long) (e.off + e.len));  // src may be *this
  return (double) (tab_add(copy.data(), copy.size()) + (off - e.off));
}
// InterpTable1D::get_telem (InterpTable1D.h:180-197) / InterpTable2D::get_xyelems (InterpTable2D.h:221-251)
int table_interval(const double* axis, double t) {
  const int n = (int) axis[0];
  const double* ts = axis + 2;
  int e;
  if (axis[1] != 0.0) {
    double q = (t - ts[0]) / (ts[1] - ts[0]);
    // the device converts with cvt.rzi.s32.f64 (saturating, NaN -> 0); clamping first gives the same interval
    if (!(q > -1.0)) q = -1.0;
    if (q > (double) n) q = (double) n;
    e = (int) q;
  } else {
    e = (int) (std::upper_bound(ts, ts + n, t) - ts) - 1;
  }
  return std::max(std::min(e, n - 2), 0);
}
int SDag::tabidx(int a, size_t axis_off) {
  if (is_const(a)) return constant((double) table_interval(&tabpool[axis_off], cval(a)));
  return raw(OP_TABIDX, a, -1, -1, (double) axis_off);
}
int SDag::tabload(int idx, size_t base_off) {
  if (is_const(idx)) return constant(tabpool.at(base_off + (size_t) (long long) cval(idx)));
  return raw(OP_TABLOAD, idx, -1, -1, (double) base_off);
}

int SDag::atan2(int y, int x) {
  if (is_const(y) && is_const(x)) return constant(std::atan2(cval(y), cval(x)));
  return raw(OP_ATAN2, y, x, -1, 0.0);
}
int SDag::cmp(uint16_t op, int a, int b) {
  if (is_const(a) && is_const(b)) {
    bool r = (op == OP_LT) ? (cval(a) < cval(b)) : (cval(a) <= cval(b));
    return constant(r ? 1.0 : 0.0);
  }
  return raw(op, a, b, -1, 0.0);
}
int SDag::select(int c, int a, int b) {
  if (is_const(c)) return (cval(c) != 0.0) ? a : b;
  if (a == b) return a;
  return raw(OP_SELECT, c, a, b, 0.0);
}

// ---------------------------------------------------------------------------------------------
// Lowering of vector nodes onto the scalar DAG
// ---------------------------------------------------------------------------------------------
static int norm_power(SDag& dag, const std::vector<int>& x, int p, int* norm_out = nullptr) {
  // ||x||^p following the reference's power cases (Norms.h:77-100): p in {1,2,3,4,-1,-2,-3,-4}
  int s2 = -1;
  for (int xi: x) {
    int sq = dag.mul(xi, xi);
    s2 = (s2 < 0) ? sq : dag.add(s2, sq);
  }
  if (p == 2) return s2;
  if (p == 4) return dag.mul(s2, s2);
  if (p == -2) return dag.unary(OP_INV, s2);
  if (p == -4) return dag.unary(OP_INV, dag.mul(s2, s2));
  int nrm = dag.unary(OP_SQRT, s2);
  if (norm_out) *norm_out = nrm;
  if (p == 1) return nrm;
  if (p == 3) return dag.mul(s2, nrm);
  if (p == -1) return dag.unary(OP_INV, nrm);
  if (p == -3) return dag.unary(OP_INV, dag.mul(s2, nrm));
  if (p == 5) return dag.mul(dag.mul(s2, s2), nrm);
  if (p == -5) return dag.unary(OP_INV, dag.mul(dag.mul(s2, s2), nrm));
  return dag.unary(OP_POWC, nrm, (double) p);
}

namespace {
struct LowerCtx {
  std::map<std::pair<int, std::vector<int>>, std::vector<int>> memo;
};
std::vector<int> lower_impl(const Graph& g, int root, const std::vector<int>& in, SDag& dag, LowerCtx& ctx);
}  // namespace

std::vector<int> lower(const Graph& g, int root, const std::vector<int>& in, SDag& dag) {
  LowerCtx ctx;
  return lower_impl(g, root, in, dag, ctx);
}

namespace {
std::vector<int> lower_impl(const Graph& g, int root, const std::vector<int>& in, SDag& dag, LowerCtx& ctx) {
  auto mkey = std::make_pair(root, in);
  auto mit = ctx.memo.find(mkey);
  if (mit != ctx.memo.end()) return mit->second;
  const VNode& nd = g.at(root);
  if ((int) in.size() != nd.ir)
    throw std::invalid_argument("lower: input size mismatch (node expects " + std::to_string(nd.ir) + ", got "
                                + std::to_string(in.size()) + ")");
  auto L = [&](int k) { return lower_impl(g, nd.ch[k], in, dag, ctx); };
  std::vector<int> out;
  switch (nd.kind) {
    case VK_ARGS: out = in; break;
    case VK_SEGMENT: {
      auto c = L(0);
      out.assign(c.begin() + nd.ip[0], c.begin() + nd.ip[0] + nd.ip[1]);
    } break;
    case VK_ELEMENTS: {
      auto c = L(0);
      for (int i: nd.ip) out.push_back(c.at(i));
    } break;
    case VK_CONSTANT:
      for (double v: nd.d) out.push_back(dag.constant(v));
      break;
    case VK_APPCONST:
      for (int i = 0; i < nd.orows; i++) out.push_back(dag.appconst(nd.ip[0] + i));
      break;
    case VK_STACK:
      for (size_t k = 0; k < nd.ch.size(); k++) {
        auto c = L((int) k);
        out.insert(out.end(), c.begin(), c.end());
      }
      break;
    case VK_SUM: {
      out = L(0);
      for (size_t k = 1; k < nd.ch.size(); k++) {
        auto c = L((int) k);
        for (size_t i = 0; i < out.size(); i++) out[i] = dag.add(out[i], c[i]);
      }
    } break;
    case VK_DIFF: {
      auto a = L(0), b = L(1);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.sub(a[i], b[i]));
    } break;
    case VK_SCALED: {
      auto a = L(0);
      for (int ai: a) out.push_back(dag.mulc(ai, nd.d[0]));
    } break;
    case VK_ROWSCALED: {
      auto a = L(0);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.mulc(a[i], nd.d[i]));
    } break;
    case VK_ADDVEC: {
      auto a = L(0);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.addc(a[i], nd.d[i]));
    } break;
    case VK_VECMINUS: {
      auto a = L(0);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.addc(dag.neg(a[i]), nd.d[i]));
    } break;
    case VK_UNARY: {
      auto a = L(0);
      uint16_t op = (uint16_t) nd.ip[0];
      double imm = nd.d.empty() ? 0.0 : nd.d[0];
      for (int ai: a) out.push_back(dag.unary(op, ai, imm));
    } break;
    case VK_NORM: {
      auto a = L(0);
      out.push_back(norm_power(dag, a, nd.ip[0]));
    } break;
    case VK_NORMALIZED: {
      auto a = L(0);
      int inv = norm_power(dag, a, -nd.ip[0]);
      for (int ai: a) out.push_back(dag.mul(ai, inv));
    } break;
    case VK_DOT: {
      auto a = L(0), b = L(1);
      int s = -1;
      for (size_t i = 0; i < a.size(); i++) {
        int p = dag.mul(a[i], b[i]);
        s = (s < 0) ? p : dag.add(s, p);
      }
      out.push_back(s);
    } break;
    case VK_CROSS: {
      auto a = L(0), b = L(1);
      out.push_back(dag.sub(dag.mul(a[1], b[2]), dag.mul(a[2], b[1])));
      out.push_back(dag.sub(dag.mul(a[2], b[0]), dag.mul(a[0], b[2])));
      out.push_back(dag.sub(dag.mul(a[0], b[1]), dag.mul(a[1], b[0])));
    } break;
    case VK_QUATPROD: {  // FunctionQuatProduct (VectorProducts.h:28-30), term order of quatprodimpl (:229-236)
      auto a = L(0), b = L(1);
      if (a.size() != 4 || b.size() != 4) throw std::invalid_argument("lower: quaternion product needs two 4-vectors");
      auto m = [&](int i, int j) { return dag.mul(a[i], b[j]); };
      out.push_back(dag.sub(dag.add(dag.add(m(0, 3), m(3, 0)), m(1, 2)), m(2, 1)));
      out.push_back(dag.sub(dag.add(dag.add(m(1, 3), m(3, 1)), m(2, 0)), m(0, 2)));
      out.push_back(dag.sub(dag.add(dag.add(m(2, 3), m(3, 2)), m(0, 1)), m(1, 0)));
      out.push_back(dag.sub(dag.sub(dag.sub(m(3, 3), m(0, 0)), m(1, 1)), m(2, 2)));
    } break;
    case VK_CWISEPROD: {
      auto a = L(0), b = L(1);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.mul(a[i], b[i]));
    } break;
    case VK_CWISEQUOT: {
      auto a = L(0), b = L(1);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.div(a[i], b[i]));
    } break;
    case VK_VSPROD: {
      auto a = L(0), s = L(1);
      for (int ai: a) out.push_back(dag.mul(ai, s[0]));
    } break;
    case VK_VSDIV: {
      auto a = L(0), s = L(1);
      for (int ai: a) out.push_back(dag.div(ai, s[0]));
    } break;
    case VK_CWISESUM: {
      auto a = L(0);
      int s = a[0];
      for (size_t i = 1; i < a.size(); i++) s = dag.add(s, a[i]);
      out.push_back(s);
    } break;
    case VK_ATAN2: {
      auto y = L(0), x = L(1);
      out.push_back(dag.atan2(y[0], x[0]));
    } break;
    case VK_NESTED: {
      auto inner = lower_impl(g, nd.ch[1], in, dag, ctx);
      out = lower_impl(g, nd.ch[0], inner, dag, ctx);
    } break;
    case VK_COMPARE: {
      auto a = L(0), b = L(1);
      switch (nd.ip[0]) {
        case 0: out.push_back(dag.cmp(OP_LT, a[0], b[0])); break;
        case 1: out.push_back(dag.cmp(OP_LE, a[0], b[0])); break;
        case 2: out.push_back(dag.cmp(OP_LT, b[0], a[0])); break;
        default: out.push_back(dag.cmp(OP_LE, b[0], a[0])); break;
      }
    } break;
    case VK_BOOL: {
      auto a = L(0), b = L(1);
      if (nd.ip[0] == 0)
        out.push_back(dag.mul(a[0], b[0]));
      else
        out.push_back(dag.select(a[0], dag.constant(1.0), b[0]));
    } break;
    case VK_IFELSE: {
      auto c = L(0), a = L(1), b = L(2);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.select(c[0], a[i], b[i]));
    } break;
    case VK_IOSCALED: {
      std::vector<int> sin(in.size());
      for (size_t i = 0; i < in.size(); i++) sin[i] = dag.mulc(in[i], nd.d[i]);
      auto a = lower_impl(g, nd.ch[0], sin, dag, ctx);
      for (size_t i = 0; i < a.size(); i++) out.push_back(dag.mulc(a[i], nd.d[nd.ir + i]));
    } break;
    case VK_INTERP1D: {  // InterpTable1D::interp_impl (InterpTable1D.h:199-262), the interpolant as arithmetic on the local
                         // coordinate; its first and second derivatives (:231-254) follow from differentiating the DAG
      auto tt = L(0);
      if (tt.size() != 1 || nd.ip.size() != 4) throw std::invalid_argument("lower: InterpTable1D takes one scalar and ip = {n, vlen, kind, even}");
      const int n = nd.ip[0], vlen = nd.ip[1], kind = nd.ip[2];
      const bool cubic = kind == 0;
      if (n < 5) throw std::invalid_argument("t coordinates must be larger than 4");
      if (vlen != nd.orows || (long long) nd.d.size() != (long long) n * (1 + vlen * (cubic ? 2 : 1)))
        throw std::invalid_argument("lower: InterpTable1D data size does not match n, vlen and the kind");
      for (int i = 0; i + 1 < n; i++)
        if (nd.d[i + 1] < nd.d[i]) throw std::invalid_argument("t Coordinates must be in ascending order");
      std::vector<double> axis(n + 2);
      axis[0] = n;
      axis[1] = nd.ip[3] ? 1.0 : 0.0;
      std::copy(nd.d.begin(), nd.d.begin() + n, axis.begin() + 2);
      const size_t aoff = dag.tab_add(axis.data(), axis.size());
      const size_t voff = dag.tab_add(nd.d.data() + n, (size_t) n * vlen);
      const size_t doff = cubic ? dag.tab_add(nd.d.data() + n + (size_t) n * vlen, (size_t) n * vlen) : 0;
      const int t = tt[0];
      const int e = dag.tabidx(t, aoff);
      const int t0 = dag.tabload(e, aoff + 2), t1 = dag.tabload(e, aoff + 3);
      const int step = dag.sub(t1, t0);
      const int u = dag.div(dag.sub(t, t0), step);
      if (cubic) {
        const int u2 = dag.mul(u, u), u3 = dag.mul(u2, u);
        // Hermite basis, InterpTable1D.h:222-225
        const int p0 = dag.addc(dag.sub(dag.mulc(u3, 2.0), dag.mulc(u2, 3.0)), 1.0);
        const int m0 = dag.mul(dag.add(dag.sub(u3, dag.mulc(u2, 2.0)), u), step);
        const int p1 = dag.add(dag.mulc(u3, -2.0), dag.mulc(u2, 3.0));
        const int m1 = dag.mul(dag.sub(u3, u2), step);
        for (int k = 0; k < vlen; k++) {
          const size_t r = (size_t) k * n;
          const int v0 = dag.tabload(e, voff + r), v1 = dag.tabload(e, voff + r + 1);
          const int d0 = dag.tabload(e, doff + r), d1 = dag.tabload(e, doff + r + 1);
          out.push_back(dag.add(dag.add(dag.add(dag.mul(v0, p0), dag.mul(v1, p1)), dag.mul(d0, m0)), dag.mul(d1, m1)));
        }
      } else {
        const int w = dag.addc(dag.neg(u), 1.0);
        for (int k = 0; k < vlen; k++) {
          const size_t r = (size_t) k * n;
          const int v0 = dag.tabload(e, voff + r), v1 = dag.tabload(e, voff + r + 1);
          out.push_back(dag.add(dag.mul(v0, w), dag.mul(v1, u)));
        }
      }
    } break;
    case VK_INTERP2D: {  // InterpTable2D::interp_impl (InterpTable2D.h:300-390): bicubic patch a = L Z R (get_amatrix, :253-298)
      auto xy = L(0);
      if (xy.size() != 2 || nd.ip.size() != 5) throw std::invalid_argument("lower: InterpTable2D takes (x, y) and ip = {nx, ny, kind, xeven, yeven}");
      const int nx = nd.ip[0], ny = nd.ip[1], kind = nd.ip[2];
      const bool cubic = kind == 0;
      if (nx < 5) throw std::invalid_argument("X coordinates must be larger than 4");
      if (ny < 5) throw std::invalid_argument("Y  coordinates must be larger than 4");
      const size_t plane = (size_t) nx * ny;
      if (nd.d.size() != (size_t) nx + ny + plane * (cubic ? 4 : 1))
        throw std::invalid_argument("lower: InterpTable2D data size does not match nx, ny and the kind");
      for (int i = 0; i + 1 < nx; i++)
        if (nd.d[i + 1] < nd.d[i]) throw std::invalid_argument("X Coordinates must be in ascending order");
      for (int i = 0; i + 1 < ny; i++)
        if (nd.d[nx + i + 1] < nd.d[nx + i]) throw std::invalid_argument("Y Coordinates must be in ascending order");
      auto axis_of = [&](const double* c, int n, int even) {
        std::vector<double> a(n + 2);
        a[0] = n;
        a[1] = even ? 1.0 : 0.0;
        std::copy(c, c + n, a.begin() + 2);
        return dag.tab_add(a.data(), a.size());
      };
      const size_t xoff = axis_of(nd.d.data(), nx, nd.ip[3]), yoff = axis_of(nd.d.data() + nx, ny, nd.ip[4]);
      size_t poff[4] = {0, 0, 0, 0};  // z, dz/dx, dz/dy, d2z/dxdy, each [ny][nx]
      for (int q = 0; q < (cubic ? 4 : 1); q++) poff[q] = dag.tab_add(nd.d.data() + nx + ny + q * plane, plane);
      const int x = xy[0], y = xy[1];
      const int ex = dag.tabidx(x, xoff), ey = dag.tabidx(y, yoff);
      const int x0 = dag.tabload(ex, xoff + 2), x1 = dag.tabload(ex, xoff + 3);
      const int y0 = dag.tabload(ey, yoff + 2), y1 = dag.tabload(ey, yoff + 3);
      const int xs = dag.sub(x1, x0), ys = dag.sub(y1, y0);
      const int xf = dag.div(dag.sub(x, x0), xs), yf = dag.div(dag.sub(y, y0), ys);
      const int cell = dag.add(dag.mulc(ey, (double) nx), ex);  // exact: indices are small integers
      auto corner = [&](int q, int dx, int dy) { return dag.tabload(cell, poff[q] + (size_t) dy * nx + dx); };
      if (cubic) {
        // Z (InterpTable2D.h:291-293): rows (z, z, dz/dx, dz/dx) at x-corner (0, 1, 0, 1); columns (.., .., d/dy, d/dy) at y-corner
        int Z[4][4];
        const int sxy = dag.mul(xs, ys);
        for (int r = 0; r < 4; r++)
          for (int c = 0; c < 4; c++) {
            const int dx = r & 1, dy = c & 1, wx = r >> 1, wy = c >> 1;
            const int q = wx ? (wy ? 3 : 1) : (wy ? 2 : 0);
            int v = corner(q, dx, dy);
            if (wx && wy)
              v = dag.mul(v, sxy);
            else if (wx)
              v = dag.mul(v, xs);
            else if (wy)
              v = dag.mul(v, ys);
            Z[r][c] = v;
          }
        static const double Lm[4][4] = {{1, 0, 0, 0}, {0, 0, 1, 0}, {-3, 3, -2, -1}, {2, -2, 1, 1}};
        static const double Rm[4][4] = {{1, 0, -3, 2}, {0, 0, 3, -2}, {0, 1, -2, 1}, {0, 0, -1, 1}};
        int LZ[4][4], Am[4][4];
        auto lin = [&](const double* w, const int* v, int stride) {
          int s = -1;
          for (int k = 0; k < 4; k++) {
            if (w[k] == 0.0) continue;
            const int term = dag.mulc(v[k * stride], w[k]);
            s = s < 0 ? term : dag.add(s, term);
          }
          return s < 0 ? dag.constant(0.0) : s;
        };
        for (int r = 0; r < 4; r++)
          for (int c = 0; c < 4; c++) LZ[r][c] = lin(Lm[r], &Z[0][c], 4);
        for (int r = 0; r < 4; r++)
          for (int c = 0; c < 4; c++) {
            const double col[4] = {Rm[0][c], Rm[1][c], Rm[2][c], Rm[3][c]};
            Am[r][c] = lin(col, &LZ[r][0], 1);
          }
        const int xf2 = dag.mul(xf, xf), yf2 = dag.mul(yf, yf);
        const int xv[4] = {dag.constant(1.0), xf, xf2, dag.mul(xf2, xf)};
        const int yv[4] = {dag.constant(1.0), yf, yf2, dag.mul(yf2, yf)};
        int z = -1;
        for (int r = 0; r < 4; r++) {
          int row = -1;
          for (int c = 0; c < 4; c++) {
            const int term = dag.mul(Am[r][c], yv[c]);
            row = row < 0 ? term : dag.add(row, term);
          }
          const int term = dag.mul(xv[r], row);
          z = z < 0 ? term : dag.add(z, term);
        }
        out.push_back(z);
      } else {  // InterpTable2D.h:363-388
        const int z00 = corner(0, 0, 0), z10 = corner(0, 1, 0), z01 = corner(0, 0, 1), z11 = corner(0, 1, 1);
        const int wx = dag.addc(dag.neg(xf), 1.0), wy = dag.addc(dag.neg(yf), 1.0);
        const int lo = dag.add(dag.mul(z00, wx), dag.mul(z10, xf)), hi = dag.add(dag.mul(z01, wx), dag.mul(z11, xf));
        out.push_back(dag.add(dag.mul(lo, wy), dag.mul(hi, yf)));
      }
    } break;
    case VK_MATPROD: {  // MatrixFunctionProduct (MatrixProduct.h:19-120): Eigen maps over the two outputs, result column-major
      auto A = L(0), B = L(1);
      if (nd.ip.size() != 5) throw std::invalid_argument("lower: matrix product takes ip = {r, k, c, A row-major, B row-major}");
      const int r = nd.ip[0], k = nd.ip[1], c = nd.ip[2];
      if ((int) A.size() != r * k || (int) B.size() != k * c || nd.orows != r * c)
        throw std::invalid_argument("Invalid matrix product. Number of columns in matrix 1 does not match number of rows in matrix 2.");
      auto a = [&](int i, int q) { return A[(size_t) (nd.ip[3] ? i * k + q : q * r + i)]; };
      auto b = [&](int q, int j) { return B[(size_t) (nd.ip[4] ? q * c + j : j * k + q)]; };
      for (int j = 0; j < c; j++)
        for (int i = 0; i < r; i++) {
          int sum = -1;
          for (int q = 0; q < k; q++) {
            const int term = dag.mul(a(i, q), b(q, j));
            sum = sum < 0 ? term : dag.add(sum, term);
          }
          out.push_back(sum);
        }
    } break;
    case VK_MATINV: {  // MatrixInverse (MatrixInverse.h:6-58); derivatives (:60-188) follow from differentiating the DAG
      auto A = L(0);
      if (nd.ip.size() != 2) throw std::invalid_argument("lower: matrix inverse takes ip = {n, row-major}");
      const int n = nd.ip[0];
      if (n < 1 || (int) A.size() != n * n) throw std::invalid_argument("Matrix must be square to be invertible");
      const bool rowmajor = nd.ip[1] != 0;
      std::vector<std::vector<int>> M((size_t) n, std::vector<int>((size_t) n)), R = M;
      for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) M[i][j] = A[(size_t) (rowmajor ? i * n + j : j * n + i)];
      if (n == 1) {
        R[0][0] = dag.unary(OP_INV, M[0][0]);
      } else if (n == 2) {  // adjugate over determinant, as Eigen does for fixed sizes up to four
        const int idet = dag.unary(OP_INV, dag.sub(dag.mul(M[0][0], M[1][1]), dag.mul(M[0][1], M[1][0])));
        R[0][0] = dag.mul(M[1][1], idet);
        R[0][1] = dag.neg(dag.mul(M[0][1], idet));
        R[1][0] = dag.neg(dag.mul(M[1][0], idet));
        R[1][1] = dag.mul(M[0][0], idet);
      } else if (n == 3) {
        auto cof = [&](int i, int j) {  // cofactor of element (i, j)
          const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
          return dag.sub(dag.mul(M[i1][j1], M[i2][j2]), dag.mul(M[i1][j2], M[i2][j1]));
        };
        int C[3][3];
        for (int i = 0; i < 3; i++)
          for (int j = 0; j < 3; j++) C[i][j] = cof(i, j);
        const int det = dag.add(dag.add(dag.mul(M[0][0], C[0][0]), dag.mul(M[0][1], C[0][1])), dag.mul(M[0][2], C[0][2]));
        const int idet = dag.unary(OP_INV, det);
        for (int i = 0; i < 3; i++)
          for (int j = 0; j < 3; j++) R[i][j] = dag.mul(C[j][i], idet);
      } else {
        // Gauss-Jordan on [M | I] with partial pivoting done by selects: before column j is eliminated, row j is exchanged
        // with every later row whose entry in that column is larger in magnitude (straight-line code; the comparisons
        // have zero derivative, so the derivatives are those of the inverse wherever the pivot order is locally constant)
        std::vector<std::vector<int>> W((size_t) n, std::vector<int>((size_t) 2 * n));
        for (int i = 0; i < n; i++)
          for (int j = 0; j < 2 * n; j++) W[i][j] = j < n ? M[i][j] : dag.constant(j - n == i ? 1.0 : 0.0);
        for (int j = 0; j < n; j++) {
          for (int i = j + 1; i < n; i++) {
            const int swap = dag.cmp(OP_LT, dag.unary(OP_ABS, W[j][j]), dag.unary(OP_ABS, W[i][j]));
            for (int q = 0; q < 2 * n; q++) {
              const int top = dag.select(swap, W[i][q], W[j][q]), bot = dag.select(swap, W[j][q], W[i][q]);
              W[j][q] = top;
              W[i][q] = bot;
            }
          }
          const int ip = dag.unary(OP_INV, W[j][j]);
          for (int q = 0; q < 2 * n; q++) W[j][q] = q == j ? dag.constant(1.0) : dag.mul(W[j][q], ip);
          for (int i = 0; i < n; i++) {
            if (i == j) continue;
            const int f = W[i][j];
            for (int q = 0; q < 2 * n; q++) W[i][q] = q == j ? dag.constant(0.0) : dag.sub(W[i][q], dag.mul(f, W[j][q]));
          }
        }
        for (int i = 0; i < n; i++)
          for (int j = 0; j < n; j++) R[i][j] = W[i][n + j];
      }
      out.resize((size_t) n * n);
      for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) out[(size_t) (rowmajor ? i * n + j : j * n + i)] = R[i][j];
    } break;
    case VK_INTERPND: {  // InterpTable3D / InterpTable4D (InterpTable3D.h:395-560, InterpTable4D.h): the tricubic / quadricubic
                         // Hermite patch through values and all mixed first derivatives at the cell corners, as a tensor product
      auto xs = L(0);
      if (nd.ip.size() < 2) throw std::invalid_argument("lower: InterpTableND takes ip = {D, kind, n_0.., even_0..}");
      const int D = nd.ip[0];
      const bool cubic = nd.ip[1] == 0;
      if (D < 1 || D > 4 || (int) xs.size() != D || (int) nd.ip.size() != 2 + 2 * D)
        throw std::invalid_argument("lower: InterpTableND dimension / coordinate count mismatch");
      std::vector<int> n((size_t) D);
      size_t cells = 1, naxis = 0;
      for (int d = 0; d < D; d++) {
        n[d] = nd.ip[2 + d];
        if (n[d] < 5) throw std::invalid_argument("coordinates must be larger than 4");
        cells *= (size_t) n[d];
        naxis += (size_t) n[d];
      }
      const int narr = cubic ? 1 << D : 1;
      if (nd.d.size() != naxis + cells * narr) throw std::invalid_argument("lower: InterpTableND data size does not match its shape");
      std::vector<size_t> stride((size_t) D);
      for (int d = D - 1, s = 1; d >= 0; d--) {
        stride[d] = (size_t) s;
        s *= n[d];
      }
      std::vector<size_t> aoff((size_t) D), arr((size_t) narr);
      size_t pos = 0;
      for (int d = 0; d < D; d++) {
        for (int i = 0; i + 1 < n[d]; i++)
          if (nd.d[pos + i + 1] < nd.d[pos + i]) throw std::invalid_argument("coordinates must be in ascending order");
        std::vector<double> axis((size_t) n[d] + 2);
        axis[0] = n[d];
        axis[1] = nd.ip[2 + D + d] ? 1.0 : 0.0;
        std::copy(nd.d.begin() + (long) pos, nd.d.begin() + (long) pos + n[d], axis.begin() + 2);
        aoff[d] = dag.tab_add(axis.data(), axis.size());
        pos += (size_t) n[d];
      }
      for (int b = 0; b < narr; b++) arr[b] = dag.tab_add(nd.d.data() + pos + (size_t) b * cells, cells);
      // per axis: interval, local coordinate, the four Hermite weights (value at 0 / 1, slope at 0 / 1 times the step)
      int cell = -1;
      std::vector<std::array<int, 4>> Bw((size_t) D);
      for (int d = 0; d < D; d++) {
        const int e = dag.tabidx(xs[d], aoff[d]);
        const int c0 = dag.tabload(e, aoff[d] + 2), c1 = dag.tabload(e, aoff[d] + 3);
        const int step = dag.sub(c1, c0);
        const int u = dag.div(dag.sub(xs[d], c0), step);
        if (cubic) {
          const int u2 = dag.mul(u, u), u3 = dag.mul(u2, u);
          Bw[d] = {dag.addc(dag.sub(dag.mulc(u3, 2.0), dag.mulc(u2, 3.0)), 1.0), dag.add(dag.mulc(u3, -2.0), dag.mulc(u2, 3.0)),
                   dag.mul(dag.add(dag.sub(u3, dag.mulc(u2, 2.0)), u), step), dag.mul(dag.sub(u3, u2), step)};
        } else {
          Bw[d] = {dag.addc(dag.neg(u), 1.0), u, -1, -1};
        }
        const int term = dag.mulc(e, (double) stride[d]);
        cell = cell < 0 ? term : dag.add(cell, term);
      }
      // V[flags][corner] -> contract one axis at a time, the last axis first
      const int per = cubic ? 2 : 1;  // flag values per axis
      std::vector<int> V;             // index: sum over axes of ((flag_d * 2 + corner_d) * 4^d)  (only used combinations filled)
      auto key = [&](const std::vector<int>& fc, int upto) {  // fc[d] = flag_d * 2 + corner_d for d < upto
        int k = 0;
        for (int d = upto - 1; d >= 0; d--) k = k * 4 + fc[d];
        return k;
      };
      V.assign((size_t) 1 << (2 * D), -1);
      std::vector<int> fc((size_t) D, 0);
      for (int fb = 0; fb < (cubic ? 1 << D : 1); fb++)
        for (int cb = 0; cb < 1 << D; cb++) {
          size_t off = 0;
          for (int d = 0; d < D; d++) {
            fc[d] = ((fb >> d & 1) << 1) | (cb >> d & 1);
            off += (size_t) (cb >> d & 1) * stride[d];
          }
          V[(size_t) key(fc, D)] = dag.tabload(cell, arr[cubic ? fb : 0] + off);
        }
      for (int d = D - 1; d >= 0; d--) {  // after this pass V is indexed by the combinations of axes 0 .. d-1
        std::vector<int> Wn((size_t) 1 << (2 * d), -1);
        for (int k = 0; k < 1 << (2 * d); k++) {
          bool used = true;  // skip combinations with a slope flag in the linear kind
          for (int q = 0, kk = k; q < d; q++, kk >>= 2)
            if ((kk & 3) >> 1 >= per) used = false;
          if (!used) continue;
          int sum = -1;
          for (int w = 0; w < 2 * per; w++) {
            const int v = V[(size_t) (k + (w << (2 * d)))];
            if (v < 0) continue;
            const int term = dag.mul(v, Bw[d][w]);
            sum = sum < 0 ? term : dag.add(sum, term);
          }
          Wn[(size_t) k] = sum;
        }
        V.swap(Wn);
      }
      out.push_back(V[0]);
    } break;
    case VK_MATVEC: {
      auto a = L(0);
      int cols = (int) a.size();
      for (int r = 0; r < nd.orows; r++) {
        int s = dag.constant(0.0);
        for (int c = 0; c < cols; c++) s = dag.add(s, dag.mulc(a[c], nd.d[(size_t) r * cols + c]));
        out.push_back(s);
      }
    } break;
    default: throw std::invalid_argument("lower: unknown node kind");
  }
  if ((int) out.size() != nd.orows) throw std::logic_error("lower: output size mismatch in kind " + std::to_string(nd.kind));
  ctx.memo.emplace(mkey, out);
  return out;
}
}  // namespace

ScalarFn lower_function(const Graph& g, int root) {
  ScalarFn f;
  f.nin = g.at(root).ir;
  f.nout = g.at(root).orows;
  std::vector<int> in(f.nin);
  for (int i = 0; i < f.nin; i++) in[i] = f.dag.input(i);
  f.outs = lower(g, root, in, f.dag);
  return f;
}

// ---------------------------------------------------------------------------------------------
// Symbolic reverse accumulation: grads[i] = d( sum_k lambda_k * f_k ) / d x_i as new DAG nodes.
// The derivative rules are the scalar content of the reference's cwise operator table
// (/root/reference/src/VectorFunctions/CommonFunctions/CwiseOperators.h:31-860) and binary products.
// ---------------------------------------------------------------------------------------------
void ScalarFn::make_gradient() {
  const int N0 = (int) dag.n.size();
  std::vector<int> adj(N0, -1);
  auto acc = [&](int i, int v) {
    if (i < 0) return;
    if (dag.is_const(v) && dag.cval(v) == 0.0) return;
    adj[i] = (adj[i] < 0) ? v : dag.add(adj[i], v);
  };
  for (int k = 0; k < nout; k++) acc(outs[k], dag.lambda(k));
  for (int i = N0 - 1; i >= 0; i--) {
    int yb = adj[i];
    if (yb < 0) continue;
    const SNode nd = dag.n[i];  // copy: dag.n may reallocate
    const int a = nd.a, b = nd.b;
    switch (nd.op) {
      case OP_CONST:
      case OP_INPUT:
      case OP_LAMBDA:
      case OP_APPCONST:
      case OP_SIGN:
      case OP_LT:
      case OP_LE:
      case OP_TABIDX:
      case OP_TABLOAD: break;
      case OP_ADD:
        acc(a, yb);
        acc(b, yb);
        break;
      case OP_SUB:
        acc(a, yb);
        acc(b, dag.neg(yb));
        break;
      case OP_MUL:
        acc(a, dag.mul(yb, b));
        acc(b, dag.mul(yb, a));
        break;
      case OP_DIV: {
        int q = dag.div(yb, b);
        acc(a, q);
        acc(b, dag.neg(dag.mul(q, i)));
      } break;
      case OP_NEG: acc(a, dag.neg(yb)); break;
      case OP_ADDC: acc(a, yb); break;
      case OP_MULC: acc(a, dag.mulc(yb, nd.imm)); break;
      case OP_SQUARE: acc(a, dag.mul(yb, dag.mulc(a, 2.0))); break;
      case OP_INV: acc(a, dag.neg(dag.mul(yb, dag.mul(i, i)))); break;
      case OP_SQRT: acc(a, dag.mul(yb, dag.mulc(dag.unary(OP_INV, i), 0.5))); break;
      case OP_EXP: acc(a, dag.mul(yb, i)); break;
      case OP_LOG: acc(a, dag.div(yb, a)); break;
      case OP_SIN: acc(a, dag.mul(yb, dag.unary(OP_COS, a))); break;
      case OP_COS: acc(a, dag.neg(dag.mul(yb, dag.unary(OP_SIN, a)))); break;
      case OP_TAN: acc(a, dag.mul(yb, dag.addc(dag.mul(i, i), 1.0))); break;
      case OP_ASIN: acc(a, dag.div(yb, dag.unary(OP_SQRT, dag.addc(dag.neg(dag.mul(a, a)), 1.0)))); break;
      case OP_ACOS: acc(a, dag.neg(dag.div(yb, dag.unary(OP_SQRT, dag.addc(dag.neg(dag.mul(a, a)), 1.0))))); break;
      case OP_ATAN: acc(a, dag.div(yb, dag.addc(dag.mul(a, a), 1.0))); break;
      case OP_SINH: acc(a, dag.mul(yb, dag.unary(OP_COSH, a))); break;
      case OP_COSH: acc(a, dag.mul(yb, dag.unary(OP_SINH, a))); break;
      case OP_TANH: acc(a, dag.mul(yb, dag.addc(dag.neg(dag.mul(i, i)), 1.0))); break;
      case OP_ASINH: acc(a, dag.div(yb, dag.unary(OP_SQRT, dag.addc(dag.mul(a, a), 1.0)))); break;
      case OP_ACOSH: acc(a, dag.div(yb, dag.unary(OP_SQRT, dag.addc(dag.mul(a, a), -1.0)))); break;
      case OP_ATANH: acc(a, dag.div(yb, dag.addc(dag.neg(dag.mul(a, a)), 1.0))); break;
      case OP_ABS: acc(a, dag.mul(yb, dag.unary(OP_SIGN, a))); break;
      case OP_POWC: acc(a, dag.mul(yb, dag.mulc(dag.unary(OP_POWC, a, nd.imm - 1.0), nd.imm))); break;
      case OP_ATAN2: {  // i = atan2(y=a, x=b)
        int r2 = dag.add(dag.mul(a, a), dag.mul(b, b));
        int q = dag.div(yb, r2);
        acc(a, dag.mul(q, b));
        acc(b, dag.neg(dag.mul(q, a)));
      } break;
      case OP_SELECT: {  // i = a ? b : c
        int zero = dag.constant(0.0);
        acc(b, dag.select(a, yb, zero));
        acc(nd.c, dag.select(a, zero, yb));
      } break;
      default: throw std::logic_error("make_gradient: unhandled op " + std::to_string(nd.op));
    }
  }
  grads.assign(nin, -1);
  for (int i = 0; i < nin; i++) {
    int id = dag.input(i);
    grads[i] = (id < N0 && adj[id] >= 0) ? adj[id] : dag.constant(0.0);
  }
}

// ---------------------------------------------------------------------------------------------
// Tape emission with liveness-based slot allocation.
// ---------------------------------------------------------------------------------------------
namespace {
struct VInstr {
  uint16_t op;
  int dstnode;  // node defined (or -1 for OUT)
  int outrow;
  int a, b, c;  // operand nodes (-1 none); for leaf ops `a` is the literal index
  double imm;
};
inline bool leaf_op(uint16_t op) { return op == OP_CONST || op == OP_INPUT || op == OP_LAMBDA || op == OP_APPCONST; }
}  // namespace

Tape emit_tape(const ScalarFn& f, bool with_adjoint) {
  const SDag& dag = f.dag;
  const int N = (int) dag.n.size();
  std::vector<char> inP(N, 0), inA(N, 0);
  auto mark = [&](int root, std::vector<char>& set) {
    std::vector<int> st {root};
    while (!st.empty()) {
      int i = st.back();
      st.pop_back();
      if (i < 0 || set[i]) continue;
      set[i] = 1;
      const SNode& nd = dag.n[i];
      if (leaf_op(nd.op)) continue;
      st.push_back(nd.a);
      if (nd.b >= 0) st.push_back(nd.b);
      if (nd.c >= 0) st.push_back(nd.c);
    }
  };
  for (int o: f.outs) mark(o, inP);
  if (with_adjoint) {
    if ((int) f.grads.size() != f.nin) throw std::logic_error("emit_tape: gradient not built");
    for (int gnode: f.grads) mark(gnode, inA);
  }
  std::vector<VInstr> vi;
  auto emit_node = [&](int i) {
    const SNode& nd = dag.n[i];
    vi.push_back(VInstr {nd.op, i, -1, nd.a, nd.b, nd.c, nd.imm});
  };
  for (int i = 0; i < N; i++)
    if (inP[i]) emit_node(i);
  for (int k = 0; k < f.nout; k++) vi.push_back(VInstr {OP_OUT_F, -1, k, f.outs[k], -1, -1, 0.0});
  const int n_primal = (int) vi.size();
  if (with_adjoint) {
    // gradient rows whose node already exists in the primal section
    for (int i = 0; i < f.nin; i++)
      if (inP[f.grads[i]]) vi.push_back(VInstr {OP_OUT_G, -1, i, f.grads[i], -1, -1, 0.0});
    for (int i = 0; i < N; i++) {
      if (!inA[i] || inP[i]) continue;
      emit_node(i);
      for (int r = 0; r < f.nin; r++)
        if (f.grads[r] == i) vi.push_back(VInstr {OP_OUT_G, -1, r, i, -1, -1, 0.0});
    }
  }
  // last use
  std::vector<int> last(N, -1);
  for (int p = 0; p < (int) vi.size(); p++) {
    const VInstr& v = vi[p];
    if (v.dstnode >= 0 && leaf_op(v.op)) continue;
    if (v.a >= 0) last[v.a] = p;
    if (v.b >= 0) last[v.b] = p;
    if (v.c >= 0) last[v.c] = p;
  }
  Tape t;
  t.nin = f.nin;
  t.nout = f.nout;
  t.has_adjoint = with_adjoint;
  t.n_primal = n_primal;
  std::vector<int> slot(N, -1);
  std::vector<int> freelist;
  int nslots = 0;
  int maxapp = -1;
  for (int p = 0; p < (int) vi.size(); p++) {
    const VInstr& v = vi[p];
    Instr ins;
    std::memset(&ins, 0, sizeof(ins));
    ins.op = v.op;
    bool leaf = (v.dstnode >= 0 && leaf_op(v.op));
    auto S = [&](int node) -> uint16_t {
      if (node < 0) return 0;
      if (slot[node] < 0) throw std::logic_error("emit_tape: operand has no slot");
      return (uint16_t) slot[node];
    };
    if (leaf) {
      ins.a = (uint16_t) std::max(v.a, 0);
      ins.imm = v.imm;
      if (v.op == OP_APPCONST) maxapp = std::max(maxapp, v.a);
    } else {
      ins.a = S(v.a);
      ins.b = S(v.b);
      if (v.op == OP_SELECT || v.op == OP_FMA)
        ins.c = S(v.c);
      else
        ins.imm = v.imm;
    }
    // release operands that die here (their slot may be reused by dst: operands are read first)
    if (!leaf) {
      int ops[3] = {v.a, v.b, v.c};
      for (int q = 0; q < 3; q++) {
        int o = ops[q];
        if (o < 0 || last[o] != p || slot[o] < 0) continue;
        bool dup = false;
        for (int r = 0; r < q; r++) dup |= (ops[r] == o);
        if (!dup) freelist.push_back(slot[o]);
      }
    }
    if (v.dstnode >= 0) {
      int s;
      if (last[v.dstnode] < 0) {
        // value never used (can happen for a dead leaf); still give it a slot and free immediately
        if (freelist.empty()) freelist.push_back(nslots++);
        s = freelist.back();
      } else if (!freelist.empty()) {
        s = freelist.back();
        freelist.pop_back();
      } else {
        s = nslots++;
      }
      slot[v.dstnode] = s;
      ins.dst = (uint16_t) s;
    } else {
      ins.dst = (uint16_t) v.outrow;
    }
    if (nslots > 65535) throw std::length_error("emit_tape: more than 65535 live slots");
    t.code.push_back(ins);
    if (p == n_primal - 1) t.slots_primal = nslots;
  }
  if (n_primal == 0) t.slots_primal = 0;
  t.slots_total = nslots;
  t.nappconst = maxapp + 1;
  return t;
}

// ---------------------------------------------------------------------------------------------
// s-expression description (consumed by the CPU oracle in tests; never by the product path)
// ---------------------------------------------------------------------------------------------
static const char* kind_name(int k) {
  static const char* names[] = {"args",     "segment", "elements",  "constant",  "stack",    "sum",     "diff",   "scaled",
                                "rowscaled", "addvec",  "vecminus",  "unary",     "norm",     "normalized", "dot", "cross",
                                "cwiseprod", "cwisequot", "vsprod",  "vsdiv",     "cwisesum", "atan2",   "nested", "ifelse",
                                "compare",  "bool",    "ioscaled",  "appconst",  "matvec",   "quatprod", "interp1d", "interp2d",
                                "matprod",  "matinv",  "interpnd"};
  return names[k];
}

std::string Graph::describe(int root) const {
  std::ostringstream os;
  os.precision(17);
  const VNode& nd = at(root);
  os << "(" << kind_name(nd.kind) << " " << nd.ir << " " << nd.orows << " " << (nd.tlin ? 1 : 0) << " [";
  for (size_t i = 0; i < nd.ip.size(); i++) os << (i ? " " : "") << nd.ip[i];
  os << "] [";
  for (size_t i = 0; i < nd.d.size(); i++) os << (i ? " " : "") << nd.d[i];
  os << "]";
  for (int c: nd.ch) os << " " << describe(c);
  os << ")";
  return os.str();
}

}  // namespace ab
