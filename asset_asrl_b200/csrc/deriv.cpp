// Symbolic second-order differentiation of a lowered function on the scalar DAG.
//
// Given f (outputs) and g = d(lambda.f)/dx (ScalarFn::make_gradient), builds DAG nodes for every Jacobian entry
// dF_k/dx_j and every lower-triangular entry of the lambda-weighted Hessian dg_i/dx_j (j <= i) by forward
// propagation of SPARSE tangent vectors: structurally zero directions are never materialised, so the scalar program
// the device runs contains only the arithmetic the sparsity of the ODE / constraint actually needs.
// The per-op rules are the scalar content of the reference's analytic derivative tables
// (/root/reference/src/VectorFunctions/CommonFunctions/CwiseOperators.h:31-860, NestedFunction.h:140-270).
#include <algorithm>
#include <stdexcept>

#include "expr.hpp"

namespace ab {

namespace {
using Tan = std::vector<std::pair<int, int>>;  // (direction, node) sorted by direction

Tan tan_scale(SDag& d, const Tan& a, int factor) {
  Tan r;
  r.reserve(a.size());
  for (auto& e: a) r.emplace_back(e.first, d.mul(e.second, factor));
  return r;
}
Tan tan_scalec(SDag& d, const Tan& a, double c) {
  Tan r;
  r.reserve(a.size());
  for (auto& e: a) r.emplace_back(e.first, d.mulc(e.second, c));
  return r;
}
// r = a + sgn * b
Tan tan_add(SDag& d, const Tan& a, const Tan& b, bool sub) {
  Tan r;
  size_t i = 0, j = 0;
  while (i < a.size() || j < b.size()) {
    if (j >= b.size() || (i < a.size() && a[i].first < b[j].first)) {
      r.push_back(a[i++]);
    } else if (i >= a.size() || b[j].first < a[i].first) {
      r.emplace_back(b[j].first, sub ? d.neg(b[j].second) : b[j].second);
      j++;
    } else {
      r.emplace_back(a[i].first, sub ? d.sub(a[i].second, b[j].second) : d.add(a[i].second, b[j].second));
      i++;
      j++;
    }
  }
  return r;
}
bool is_zero(const SDag& d, int n) { return d.is_const(n) && d.cval(n) == 0.0; }
void prune(const SDag& d, Tan& t) {
  t.erase(std::remove_if(t.begin(), t.end(), [&](const std::pair<int, int>& e) { return is_zero(d, e.second); }), t.end());
}
}  // namespace

void ScalarFn::make_second_order() {
  if ((int) grads.size() != nin) make_gradient();
  SDag& d = dag;
  const int N0 = (int) d.n.size();
  // nodes needed: reachable from outs and grads
  std::vector<char> need(N0, 0);
  {
    std::vector<int> st(outs.begin(), outs.end());
    st.insert(st.end(), grads.begin(), grads.end());
    while (!st.empty()) {
      int i = st.back();
      st.pop_back();
      if (i < 0 || need[i]) continue;
      need[i] = 1;
      const SNode& nd = d.n[i];
      if (nd.op == OP_CONST || nd.op == OP_INPUT || nd.op == OP_LAMBDA || nd.op == OP_APPCONST) continue;
      st.push_back(nd.a);
      if (nd.b >= 0) st.push_back(nd.b);
      if (nd.c >= 0) st.push_back(nd.c);
    }
  }
  std::vector<Tan> T(N0);
  const int one = d.constant(1.0);
  for (int i = 0; i < N0; i++) {
    if (!need[i]) continue;
    const SNode nd = d.n[i];  // copy: the DAG grows below
    const int a = nd.a, b = nd.b;
    Tan t;
    auto unary = [&](int fprime) { t = tan_scale(d, T[a], fprime); };
    switch (nd.op) {
      case OP_CONST:
      case OP_LAMBDA:
      case OP_APPCONST:
      case OP_SIGN:
      case OP_LT:
      case OP_LE:
      case OP_TABIDX:
      case OP_TABLOAD: break;
      case OP_INPUT: t.emplace_back(a, one); break;
      case OP_ADD: t = tan_add(d, T[a], T[b], false); break;
      case OP_SUB: t = tan_add(d, T[a], T[b], true); break;
      case OP_MUL: t = tan_add(d, tan_scale(d, T[a], b), tan_scale(d, T[b], a), false); break;
      case OP_DIV: {  // i = a / b: (a' - i b') / b
        if (T[a].empty() && T[b].empty()) break;
        const int q = d.unary(OP_INV, b);
        t = tan_scale(d, tan_add(d, T[a], tan_scale(d, T[b], i), true), q);
      } break;
      case OP_NEG: t = tan_scalec(d, T[a], -1.0); break;
      case OP_ADDC: t = T[a]; break;
      case OP_MULC: t = tan_scalec(d, T[a], nd.imm); break;
      case OP_SQUARE: if (!T[a].empty()) unary(d.mulc(a, 2.0)); break;
      case OP_INV: if (!T[a].empty()) unary(d.neg(d.mul(i, i))); break;
      case OP_SQRT: if (!T[a].empty()) unary(d.mulc(d.unary(OP_INV, i), 0.5)); break;
      case OP_EXP: if (!T[a].empty()) unary(i); break;
      case OP_LOG: if (!T[a].empty()) unary(d.unary(OP_INV, a)); break;
      case OP_SIN: if (!T[a].empty()) unary(d.unary(OP_COS, a)); break;
      case OP_COS: if (!T[a].empty()) unary(d.neg(d.unary(OP_SIN, a))); break;
      case OP_TAN: if (!T[a].empty()) unary(d.addc(d.mul(i, i), 1.0)); break;
      case OP_ASIN: if (!T[a].empty()) unary(d.unary(OP_INV, d.unary(OP_SQRT, d.addc(d.neg(d.mul(a, a)), 1.0)))); break;
      case OP_ACOS: if (!T[a].empty()) unary(d.neg(d.unary(OP_INV, d.unary(OP_SQRT, d.addc(d.neg(d.mul(a, a)), 1.0))))); break;
      case OP_ATAN: if (!T[a].empty()) unary(d.unary(OP_INV, d.addc(d.mul(a, a), 1.0))); break;
      case OP_SINH: if (!T[a].empty()) unary(d.unary(OP_COSH, a)); break;
      case OP_COSH: if (!T[a].empty()) unary(d.unary(OP_SINH, a)); break;
      case OP_TANH: if (!T[a].empty()) unary(d.addc(d.neg(d.mul(i, i)), 1.0)); break;
      case OP_ASINH: if (!T[a].empty()) unary(d.unary(OP_INV, d.unary(OP_SQRT, d.addc(d.mul(a, a), 1.0)))); break;
      case OP_ACOSH: if (!T[a].empty()) unary(d.unary(OP_INV, d.unary(OP_SQRT, d.addc(d.mul(a, a), -1.0)))); break;
      case OP_ATANH: if (!T[a].empty()) unary(d.unary(OP_INV, d.addc(d.neg(d.mul(a, a)), 1.0))); break;
      case OP_ABS: if (!T[a].empty()) unary(d.unary(OP_SIGN, a)); break;
      case OP_POWC: if (!T[a].empty()) unary(d.mulc(d.unary(OP_POWC, a, nd.imm - 1.0), nd.imm)); break;
      case OP_ATAN2: {  // i = atan2(y = a, x = b): (x y' - y x') / (x^2 + y^2)
        if (T[a].empty() && T[b].empty()) break;
        const int q = d.unary(OP_INV, d.add(d.mul(a, a), d.mul(b, b)));
        t = tan_add(d, tan_scale(d, T[a], d.mul(b, q)), tan_scale(d, T[b], d.mul(a, q)), true);
      } break;
      case OP_SELECT: {  // i = a ? b : c
        const Tan &tb = T[b], &tc = T[nd.c];
        const int zero = d.constant(0.0);
        size_t p = 0, q = 0;
        while (p < tb.size() || q < tc.size()) {
          if (q >= tc.size() || (p < tb.size() && tb[p].first < tc[q].first)) {
            t.emplace_back(tb[p].first, d.select(a, tb[p].second, zero));
            p++;
          } else if (p >= tb.size() || tc[q].first < tb[p].first) {
            t.emplace_back(tc[q].first, d.select(a, zero, tc[q].second));
            q++;
          } else {
            t.emplace_back(tb[p].first, d.select(a, tb[p].second, tc[q].second));
            p++;
            q++;
          }
        }
      } break;
      case OP_FMA: {  // a*b + c
        t = tan_add(d, tan_add(d, tan_scale(d, T[a], b), tan_scale(d, T[b], a), false), T[nd.c], false);
      } break;
      default: throw std::logic_error("make_second_order: unhandled op " + std::to_string(nd.op));
    }
    prune(d, t);
    T[i] = std::move(t);
  }
  const int zero = d.constant(0.0);
  jac.assign((size_t) nout * nin, zero);
  for (int k = 0; k < nout; k++)
    for (auto& e: T[outs[k]]) jac[(size_t) k * nin + e.first] = e.second;
  hess.assign((size_t) nin * nin, zero);  // full square, only j <= i is used by the emitter
  for (int i = 0; i < nin; i++) {
    const int gi = grads[i];
    if (gi >= N0) continue;  // a constant created after the sweep (zero gradient)
    for (auto& e: T[gi]) hess[(size_t) i * nin + e.first] = e.second;
  }
}

}  // namespace ab
