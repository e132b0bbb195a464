// oracle.cpp -- CPU restatement of ASSET's KKT evaluation hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the checker for the CUDA path in asset_asrl_b200/: only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load it.  The product never links or calls it.
//
// PARITY STATUS: "parity unpinned" at the reference-binary level.  The reference (AlabamaASRL/asset_asrl
// @ 8d51f95) cannot be compiled in this image (its Eigen fork, autodiff submodule and Intel MKL are absent,
// SURVEY.md section 8c) and its tests hold no per-iterate golden vectors.  The restatement below follows the
// reference's algorithm function by function (citations on each routine); it is pinned transitively by
//   (1) the reference's own derivative self-checks (jx^T L == gx to 1e-12, finite differences to 1e-4,
//       asset_asrl/test/test_VectorFunctions/__init__.py:11-67), applied in tests/test_oracle.py,
//   (2) independent 50-digit mpmath evaluations of the same formulas committed under tests/golden/.
//
// What is restated, in the reference's own evaluation style (dense analytic derivatives, NOT the product's
// scalar tape / forward-over-reverse scheme, so the two paths share no arithmetic):
//   * expression nodes: value, dense Jacobian, adjoint-weighted Hessian via the NestedFunction chain rule
//       J = Jo*Ji,  H = Ji^T Ho Ji + Hi(lambda <- Jo^T lambda)      (src/VectorFunctions/CommonFunctions/NestedFunction.h:140-270)
//   * LGLDefects<CS> with DI_DCS / C_AVS / HTpar                   (src/OptimalControl/LGLDefects.h:289-551)
//   * TrapezoidalDefects                                           (src/OptimalControl/TrapezoidalDefects.h:263-435)
//   * LGLIntegral, LGLControlSpline, LGLMeshSpacing, SingleMeshSpacing
//   * getKKTSpace triplet order, analyzeSparsity (upper-triangular CSR + KKTLocations)
//       (src/VectorFunctions/DenseFunctionBase.h:1097-1129, src/Solvers/NonLinearProgram.cpp:267-344)
//   * evalKKT's gather -> compute -> KKTFillAll scatter, RHS fill, thread split by contiguous application chunks
//       (src/Solvers/NonLinearProgram.cpp:473-537, DenseFunctionBase.h:1276-1466, IndexingData.h:117-146)
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

namespace orc {

// ------------------------------------------------------------------------------------------------
// tiny dense helpers (row-major)
// ------------------------------------------------------------------------------------------------
struct Mat {
  int r = 0, c = 0;
  std::vector<double> a;
  Mat() {}
  Mat(int r_, int c_) : r(r_), c(c_), a((size_t) r_ * c_, 0.0) {}
  void reset(int r_, int c_) {
    r = r_;
    c = c_;
    a.assign((size_t) r_ * c_, 0.0);
  }
  double& operator()(int i, int j) { return a[(size_t) i * c + j]; }
  double operator()(int i, int j) const { return a[(size_t) i * c + j]; }
};
using Vec = std::vector<double>;

// C += alpha * A * B
static void gemm_acc(Mat& C, const Mat& A, const Mat& B, double alpha = 1.0) {
  for (int i = 0; i < A.r; i++)
    for (int k = 0; k < A.c; k++) {
      const double aik = alpha * A(i, k);
      if (aik == 0.0) continue;
      const double* b = &B.a[(size_t) k * B.c];
      double* c = &C.a[(size_t) i * C.c];
      for (int j = 0; j < B.c; j++) c[j] += aik * b[j];
    }
}
// C += alpha * A^T * M * A   (M is k x k, A is k x n)
static void congruence_acc(Mat& C, const Mat& A, const Mat& M, double alpha = 1.0) {
  Mat T(M.r, A.c);
  gemm_acc(T, M, A);
  for (int k = 0; k < A.r; k++)
    for (int i = 0; i < A.c; i++) {
      const double aki = alpha * A(k, i);
      if (aki == 0.0) continue;
      const double* t = &T.a[(size_t) k * T.c];
      double* c = &C.a[(size_t) i * C.c];
      for (int j = 0; j < A.c; j++) c[j] += aki * t[j];
    }
}
// y = A^T * l
static Vec atv(const Mat& A, const double* l) {
  Vec y(A.c, 0.0);
  for (int i = 0; i < A.r; i++)
    for (int j = 0; j < A.c; j++) y[j] += A(i, j) * l[i];
  return y;
}

// ------------------------------------------------------------------------------------------------
// expression tree (parsed from the s-expression the product's Graph::describe emits)
// ------------------------------------------------------------------------------------------------
enum Kind {
  K_ARGS, K_SEGMENT, K_ELEMENTS, K_CONSTANT, K_STACK, K_SUM, K_DIFF, K_SCALED, K_ROWSCALED, K_ADDVEC, K_VECMINUS,
  K_UNARY, K_NORM, K_NORMALIZED, K_DOT, K_CROSS, K_CWISEPROD, K_CWISEQUOT, K_VSPROD, K_VSDIV, K_CWISESUM, K_ATAN2,
  K_NESTED, K_IFELSE, K_COMPARE, K_BOOL, K_IOSCALED, K_APPCONST, K_MATVEC, K_QUATPROD, K_INTERP1D, K_INTERP2D, K_MATPROD, K_MATINV, K_INTERPND, K_COUNT
};
static const char* kKindNames[K_COUNT] = {"args", "segment", "elements", "constant", "stack", "sum", "diff", "scaled",
                                          "rowscaled", "addvec", "vecminus", "unary", "norm", "normalized", "dot", "cross",
                                          "cwiseprod", "cwisequot", "vsprod", "vsdiv", "cwisesum", "atan2", "nested", "ifelse",
                                          "compare", "bool", "ioscaled", "appconst", "matvec", "quatprod", "interp1d", "interp2d", "matprod", "matinv", "interpnd"};

// unary opcodes: numeric values of ab::Op in asset_asrl_b200/csrc/tape.h (the describe() string carries them)
enum UOp { U_NEG = 8, U_SQUARE = 11, U_INV, U_SQRT, U_EXP, U_LOG, U_SIN, U_COS, U_TAN, U_ASIN, U_ACOS, U_ATAN, U_SINH, U_COSH,
           U_TANH, U_ASINH, U_ACOSH, U_ATANH, U_ABS, U_SIGN, U_POWC };

struct Node {
  int kind = 0, ir = 0, orows = 0;
  std::vector<int> ip;
  std::vector<double> d;
  std::vector<std::unique_ptr<Node>> ch;
  // interpolation tables: the stencil derivatives of the table data, computed on first use (they depend on the data only)
  mutable std::vector<std::vector<double>> aux;
  mutable std::once_flag aux_once;
};

struct Parser {
  const std::string& s;
  size_t p = 0;
  explicit Parser(const std::string& s_) : s(s_) {}
  void ws() {
    while (p < s.size() && isspace((unsigned char) s[p])) p++;
  }
  void expect(char c) {
    ws();
    if (p >= s.size() || s[p] != c) throw std::invalid_argument(std::string("oracle parse: expected '") + c + "' at " + std::to_string(p));
    p++;
  }
  std::string word() {
    ws();
    size_t b = p;
    while (p < s.size() && !isspace((unsigned char) s[p]) && s[p] != '(' && s[p] != ')' && s[p] != '[' && s[p] != ']') p++;
    return s.substr(b, p - b);
  }
  std::unique_ptr<Node> node() {
    expect('(');
    auto n = std::make_unique<Node>();
    std::string k = word();
    n->kind = -1;
    for (int i = 0; i < K_COUNT; i++)
      if (k == kKindNames[i]) n->kind = i;
    if (n->kind < 0) throw std::invalid_argument("oracle parse: unknown kind " + k);
    n->ir = std::stoi(word());
    n->orows = std::stoi(word());
    word();  // type-level linearity flag: not needed for arithmetic
    expect('[');
    for (ws(); s[p] != ']'; ws()) n->ip.push_back(std::stoi(word()));
    expect(']');
    expect('[');
    for (ws(); s[p] != ']'; ws()) n->d.push_back(std::stod(word()));
    expect(']');
    for (ws(); p < s.size() && s[p] == '('; ws()) n->ch.push_back(node());
    expect(')');
    return n;
  }
};

// scalar op table: f, f', f''.  Restates CwiseOperators.h:31-860 (per-op derivative formulas).
static void unary_eval(int op, double imm, double x, double& f, double& d1, double& d2) {
  switch (op) {
    case U_NEG: f = -x; d1 = -1; d2 = 0; break;
    case U_SQUARE: f = x * x; d1 = 2 * x; d2 = 2; break;
    case U_INV: f = 1 / x; d1 = -1 / (x * x); d2 = 2 / (x * x * x); break;
    case U_SQRT: f = std::sqrt(x); d1 = 0.5 / f; d2 = -0.5 * d1 / x; break;
    case U_EXP: f = std::exp(x); d1 = f; d2 = f; break;
    case U_LOG: f = std::log(x); d1 = 1 / x; d2 = -1 / (x * x); break;
    case U_SIN: f = std::sin(x); d1 = std::cos(x); d2 = -f; break;
    case U_COS: f = std::cos(x); d1 = -std::sin(x); d2 = -f; break;
    case U_TAN: {
      f = std::tan(x);
      const double c = std::cos(x);
      d1 = 1 / (c * c);
      d2 = 2 * f * d1;
    } break;
    case U_ASIN: f = std::asin(x); d1 = 1 / std::sqrt(1 - x * x); d2 = x * d1 * d1 * d1; break;
    case U_ACOS: f = std::acos(x); d1 = -1 / std::sqrt(1 - x * x); d2 = x * d1 * d1 * d1; break;
    case U_ATAN: f = std::atan(x); d1 = 1 / (1 + x * x); d2 = -2 * x * d1 * d1; break;
    case U_SINH: f = std::sinh(x); d1 = std::cosh(x); d2 = f; break;
    case U_COSH: f = std::cosh(x); d1 = std::sinh(x); d2 = f; break;
    case U_TANH: f = std::tanh(x); d1 = 1 - f * f; d2 = -2 * f * d1; break;
    case U_ASINH: f = std::asinh(x); d1 = 1 / std::sqrt(x * x + 1); d2 = -x * d1 * d1 * d1; break;
    case U_ACOSH: f = std::acosh(x); d1 = 1 / std::sqrt(x * x - 1); d2 = -x * d1 * d1 * d1; break;
    case U_ATANH: f = std::atanh(x); d1 = 1 / (1 - x * x); d2 = 2 * x * d1 * d1; break;
    case U_ABS: f = std::fabs(x); d1 = (x > 0) - (x < 0); d2 = 0; break;
    case U_SIGN: f = (x > 0) - (x < 0); d1 = 0; d2 = 0; break;
    case U_POWC: f = std::pow(x, imm); d1 = imm * std::pow(x, imm - 1); d2 = imm * (imm - 1) * std::pow(x, imm - 2); break;
    default: throw std::invalid_argument("oracle: unknown unary op " + std::to_string(op));
  }
}

// ------------------------------------------------------------------------------------------------
// interpolation tables: InterpTable1D.h:9-262, InterpTable2D.h:7-390 restated (own stencil derivatives, own interval
// search, the reference's closed-form first and second derivatives of the interpolants)
// ------------------------------------------------------------------------------------------------
// weights of the five-point stencil for d/dc at c[i] over samples c[start .. start+5): solves sum_k w_k tau_k^r = [r == 1]
static void stencil5(const double* c, int n, int i, int& start, double w[5]) {
  if (i + 2 <= n - 1 && i - 2 >= 0)
    start = i - 2;
  else if (i < n - 1 - i)
    start = 0;
  else
    start = n - 5;
  const double step = std::fabs(c[i + (i < n - 1 ? 1 : -1)] - c[i]);
  long double A[5][6];
  for (int k = 0; k < 5; k++) {
    const long double tau = ((long double) c[start + k] - c[i]) / step;
    long double pw = 1;
    for (int r = 0; r < 5; r++) {
      A[r][k] = pw;
      pw *= tau;
    }
  }
  for (int r = 0; r < 5; r++) A[r][5] = (r == 1);
  for (int col = 0; col < 5; col++) {  // Gauss-Jordan with partial pivoting
    int piv = col;
    for (int r = col + 1; r < 5; r++)
      if (fabsl(A[r][col]) > fabsl(A[piv][col])) piv = r;
    for (int k = 0; k < 6; k++) std::swap(A[col][k], A[piv][k]);
    for (int r = 0; r < 5; r++) {
      if (r == col) continue;
      const long double f = A[r][col] / A[col][col];
      for (int k = col; k < 6; k++) A[r][k] -= f * A[col][k];
    }
  }
  for (int k = 0; k < 5; k++) w[k] = (double) (A[k][5] / A[k][k] / step);
}
// derivative along a strided axis: out[j*stride] = d/dc of v[j*stride], j < n
static void stencil_axis(const double* c, int n, const double* v, double* out, size_t stride) {
  for (int i = 0; i < n; i++) {
    int start;
    double w[5];
    stencil5(c, n, i, start, w);
    double s = 0;
    for (int k = 0; k < 5; k++) s += w[k] * v[(size_t) (start + k) * stride];
    out[(size_t) i * stride] = s;
  }
}
static int interval_of(const double* c, int n, bool even, double t) {
  int e;
  if (even) {
    const double q = (t - c[0]) / (c[1] - c[0]);
    e = q < 0 ? 0 : (q > n ? n : (int) q);
  } else {
    e = (int) (std::upper_bound(c, c + n, t) - c) - 1;
  }
  return std::max(std::min(e, n - 2), 0);
}

struct Eval {
  const double* appc = nullptr;
  // fx = f(x); if J: J (orows x ir) assigned; if H: H (ir x ir) += Hessian of l.f
  void run(const Node& n, const double* x, double* fx, Mat* J, const double* l, Mat* H) const;
  // outer function phi on the stacked child outputs Y (size ny): value, Jphi (or x ny), Hphi(l) (ny x ny)
  void outer(const Node& n, const Vec& Y, const std::vector<int>& off, double* fx, Mat* Jp, const double* l, Mat* Hp) const;
};

void Eval::outer(const Node& n, const Vec& Y, const std::vector<int>& off, double* fx, Mat* Jp, const double* l, Mat* Hp) const {
  const int ny = (int) Y.size();
  const int orr = n.orows;
  switch (n.kind) {
    case K_SEGMENT:  // Segment.h: rows [start, start+size) of the child
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[n.ip[0] + i];
        if (Jp) (*Jp)(i, n.ip[0] + i) = 1.0;
      }
      break;
    case K_ELEMENTS:  // Elements.h
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[n.ip[i]];
        if (Jp) (*Jp)(i, n.ip[i]) = 1.0;
      }
      break;
    case K_STACK:  // Stacked.h:74-454
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[i];
        if (Jp) (*Jp)(i, i) = 1.0;
      }
      break;
    case K_SUM:  // Summation.h:106-614
      for (int i = 0; i < orr; i++) {
        fx[i] = 0;
        for (size_t k = 0; k < n.ch.size(); k++) {
          fx[i] += Y[off[k] + i];
          if (Jp) (*Jp)(i, off[k] + i) = 1.0;
        }
      }
      break;
    case K_DIFF:
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[i] - Y[off[1] + i];
        if (Jp) {
          (*Jp)(i, i) = 1.0;
          (*Jp)(i, off[1] + i) = -1.0;
        }
      }
      break;
    case K_SCALED:  // Scaled.h:74-637
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[i] * n.d[0];
        if (Jp) (*Jp)(i, i) = n.d[0];
      }
      break;
    case K_ROWSCALED:
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[i] * n.d[i];
        if (Jp) (*Jp)(i, i) = n.d[i];
      }
      break;
    case K_ADDVEC:  // FunctionVectorSums.h
      for (int i = 0; i < orr; i++) {
        fx[i] = Y[i] + n.d[i];
        if (Jp) (*Jp)(i, i) = 1.0;
      }
      break;
    case K_VECMINUS:
      for (int i = 0; i < orr; i++) {
        fx[i] = n.d[i] - Y[i];
        if (Jp) (*Jp)(i, i) = -1.0;
      }
      break;
    case K_MATVEC:
      for (int i = 0; i < orr; i++) {
        fx[i] = 0;
        for (int j = 0; j < ny; j++) {
          fx[i] += n.d[(size_t) i * ny + j] * Y[j];
          if (Jp) (*Jp)(i, j) = n.d[(size_t) i * ny + j];
        }
      }
      break;
    case K_UNARY:  // CwiseFunctionOperator, CwiseOperators.h:872-960
      for (int i = 0; i < orr; i++) {
        double f, d1, d2;
        unary_eval(n.ip[0], n.d.empty() ? 0.0 : n.d[0], Y[i], f, d1, d2);
        fx[i] = f;
        if (Jp) (*Jp)(i, i) = d1;
        if (Hp) (*Hp)(i, i) += l[0 + i] * d2;
      }
      break;
    case K_NORM: {  // Norms.h:48-192: ||y||^p
      const int p = n.ip[0];
      double s2 = 0;
      for (double v: Y) s2 += v * v;
      const double r = std::sqrt(s2);
      fx[0] = std::pow(r, (double) p);
      if (Jp || Hp) {
        const double g = p * std::pow(r, (double) (p - 2));             // grad = g * y
        const double h = p * (p - 2) * std::pow(r, (double) (p - 4));   // hess = g I + h y y^T
        for (int i = 0; i < ny; i++) {
          if (Jp) (*Jp)(0, i) = g * Y[i];
          if (Hp) {
            (*Hp)(i, i) += l[0] * g;
            for (int j = 0; j < ny; j++) (*Hp)(i, j) += l[0] * h * Y[i] * Y[j];
          }
        }
      }
    } break;
    case K_NORMALIZED: {  // Normalized.h:27-199: y / ||y||^p
      const int p = n.ip[0];
      double s2 = 0;
      for (double v: Y) s2 += v * v;
      const double r = std::sqrt(s2);
      const double a = std::pow(r, (double) -p), b = p * std::pow(r, (double) (-p - 2)),
                   c = p * (p + 2) * std::pow(r, (double) (-p - 4));
      double ly = 0;
      for (int i = 0; i < ny; i++) {
        fx[i] = Y[i] * a;
        if (Hp) ly += l[i] * Y[i];
      }
      for (int i = 0; i < ny; i++)
        for (int j = 0; j < ny; j++) {
          if (Jp) (*Jp)(i, j) = (i == j ? a : 0.0) - b * Y[i] * Y[j];
          // sum_i l_i d2phi_i/dy_j dy_k = -b (l_j y_k + l_k y_j + delta_jk l.y) + c (l.y) y_j y_k
          if (Hp) (*Hp)(i, j) += -b * (l[i] * Y[j] + l[j] * Y[i] + (i == j ? ly : 0.0)) + c * ly * Y[i] * Y[j];
        }
    } break;
    case K_DOT: {  // DotProduct.h
      const int m = off[1];
      fx[0] = 0;
      for (int i = 0; i < m; i++) {
        fx[0] += Y[i] * Y[m + i];
        if (Jp) {
          (*Jp)(0, i) = Y[m + i];
          (*Jp)(0, m + i) = Y[i];
        }
        if (Hp) {
          (*Hp)(i, m + i) += l[0];
          (*Hp)(m + i, i) += l[0];
        }
      }
    } break;
    case K_CROSS: {  // VectorProducts.h
      const double* a = &Y[0];
      const double* b = &Y[3];
      fx[0] = a[1] * b[2] - a[2] * b[1];
      fx[1] = a[2] * b[0] - a[0] * b[2];
      fx[2] = a[0] * b[1] - a[1] * b[0];
      static const int nx[3] = {1, 2, 0}, pv[3] = {2, 0, 1};
      for (int i = 0; i < 3; i++) {
        const int j = nx[i], k = pv[i];  // f_i = a_j b_k - a_k b_j
        if (Jp) {
          (*Jp)(i, j) = b[k];
          (*Jp)(i, k) = -b[j];
          (*Jp)(i, 3 + k) = a[j];
          (*Jp)(i, 3 + j) = -a[k];
        }
        if (Hp) {
          (*Hp)(j, 3 + k) += l[i];
          (*Hp)(3 + k, j) += l[i];
          (*Hp)(k, 3 + j) -= l[i];
          (*Hp)(3 + j, k) -= l[i];
        }
      }
    } break;
    case K_QUATPROD: {  // VectorProducts.h:28-30 (FunctionQuatProduct), :229-236 (quatprodimpl), :238-262 (fillquatprodmatrix)
      // out_o = sum of sg * a_i * b_j over the bilinear terms below (quaternions stored x, y, z, w)
      static const int T[16][4] = {{0, 0, 3, 1}, {0, 3, 0, 1}, {0, 1, 2, 1}, {0, 2, 1, -1}, {1, 1, 3, 1}, {1, 3, 1, 1}, {1, 2, 0, 1}, {1, 0, 2, -1},
                                   {2, 2, 3, 1}, {2, 3, 2, 1}, {2, 0, 1, 1}, {2, 1, 0, -1}, {3, 3, 3, 1}, {3, 0, 0, -1}, {3, 1, 1, -1}, {3, 2, 2, -1}};
      const double* a = &Y[0];
      const double* b = &Y[4];
      for (int o = 0; o < 4; o++) fx[o] = 0.0;
      for (const auto& t: T) {
        const int o = t[0], i = t[1], j = t[2];
        const double sg = t[3];
        fx[o] += sg * a[i] * b[j];
        if (Jp) {
          (*Jp)(o, i) += sg * b[j];
          (*Jp)(o, 4 + j) += sg * a[i];
        }
        if (Hp) {
          (*Hp)(i, 4 + j) += sg * l[o];
          (*Hp)(4 + j, i) += sg * l[o];
        }
      }
    } break;
    case K_INTERP1D: {  // InterpFunction1D (InterpTable1D.h:322-400) over interp_impl (:199-262)
      const int nt = n.ip[0], vlen = n.ip[1];
      const bool cubic = n.ip[2] == 0, even = n.ip[3] != 0;
      const double* ts = n.d.data();
      const double* vs = ts + nt;  // [vlen][nt]
      const double t = Y[0];
      const int e = interval_of(ts, nt, even, t);
      const double h = ts[e + 1] - ts[e], u = (t - ts[e]) / h;
      if (cubic)  // slopes recomputed here (calc_derivs, :133-178), not taken from the description
        std::call_once(n.aux_once, [&]() {
          n.aux.assign(1, Vec((size_t) vlen * nt));
          for (int k = 0; k < vlen; k++) stencil_axis(ts, nt, vs + (size_t) k * nt, &n.aux[0][(size_t) k * nt], 1);
        });
      const Vec& dv = cubic ? n.aux[0] : n.d;
      double h2 = 0;
      for (int k = 0; k < vlen; k++) {
        const double v0 = vs[(size_t) k * nt + e], v1 = vs[(size_t) k * nt + e + 1];
        double f, d1, d2;
        if (cubic) {
          const double s0 = dv[(size_t) k * nt + e], s1 = dv[(size_t) k * nt + e + 1], u2 = u * u, u3 = u2 * u;
          f = v0 * (2 * u3 - 3 * u2 + 1) + v1 * (-2 * u3 + 3 * u2) + s0 * ((u3 - 2 * u2 + u) * h) + s1 * ((u3 - u2) * h);
          d1 = v0 * ((6 * u2 - 6 * u) / h) + v1 * ((-6 * u2 + 6 * u) / h) + s0 * (3 * u2 - 4 * u + 1) + s1 * (3 * u2 - 2 * u);
          d2 = v0 * ((12 * u - 6) / (h * h)) + v1 * ((-12 * u + 6) / (h * h)) + s0 * ((6 * u - 4) / h) + s1 * ((6 * u - 2) / h);
        } else {
          f = v0 * (1 - u) + v1 * u;
          d1 = (v1 - v0) / h;
          d2 = 0;
        }
        fx[k] = f;
        if (Jp) (*Jp)(k, 0) = d1;
        if (Hp) h2 += l[k] * d2;
      }
      if (Hp) (*Hp)(0, 0) += h2;
    } break;
    case K_INTERP2D: {  // InterpFunction2D (InterpTable2D.h:428-485) over interp_impl (:300-390)
      const int nx = n.ip[0], ny = n.ip[1];
      const bool cubic = n.ip[2] == 0;
      const double* xs = n.d.data();
      const double* ys = xs + nx;
      const double* zs = ys + ny;  // [ny][nx]
      const double x = Y[0], y = Y[1];
      const int ex = interval_of(xs, nx, n.ip[3] != 0, x), ey = interval_of(ys, ny, n.ip[4] != 0, y);
      const double hx = xs[ex + 1] - xs[ex], hy = ys[ey + 1] - ys[ey];
      const double xf = (x - xs[ex]) / hx, yf = (y - ys[ey]) / hy;
      auto at = [&](const double* p, int dx, int dy) { return p[(size_t) (ey + dy) * nx + ex + dx]; };
      double z, zx, zy, zxx, zxy, zyy;
      if (cubic) {
        const size_t plane = (size_t) nx * ny;
        std::call_once(n.aux_once, [&]() {  // calc_derivs (:122-211)
          n.aux.assign(3, Vec(plane));
          Vec &ax_ = n.aux[0], &ay_ = n.aux[1], &axy_ = n.aux[2];
          for (int c = 0; c < nx; c++) stencil_axis(ys, ny, zs + c, &ay_[c], nx);
          for (int r = 0; r < ny; r++) {
            stencil_axis(xs, nx, zs + (size_t) r * nx, &ax_[(size_t) r * nx], 1);
            stencil_axis(xs, nx, &ay_[(size_t) r * nx], &axy_[(size_t) r * nx], 1);
          }
        });
        const Vec &dzx = n.aux[0], &dzy = n.aux[1], &dzxy = n.aux[2];
        // Hermite patch in tensor form: sum over corners of value / x-slope / y-slope / twist times 1-D basis products;
        // the same polynomial as xvec^T (L Z R) yvec of get_amatrix (:253-298)
        auto basis = [](double s, double hh, double B[4], double D1[4], double D2[4]) {
          const double s2 = s * s, s3 = s2 * s;  // value at 0, value at 1, slope at 0, slope at 1 (slopes scaled by the step)
          B[0] = 2 * s3 - 3 * s2 + 1, B[1] = -2 * s3 + 3 * s2, B[2] = (s3 - 2 * s2 + s) * hh, B[3] = (s3 - s2) * hh;
          D1[0] = (6 * s2 - 6 * s) / hh, D1[1] = -D1[0], D1[2] = 3 * s2 - 4 * s + 1, D1[3] = 3 * s2 - 2 * s;
          D2[0] = (12 * s - 6) / (hh * hh), D2[1] = -D2[0], D2[2] = (6 * s - 4) / hh, D2[3] = (6 * s - 2) / hh;
        };
        double Bx[4], Dx[4], DDx[4], By[4], Dy[4], DDy[4];
        basis(xf, hx, Bx, Dx, DDx);
        basis(yf, hy, By, Dy, DDy);
        z = zx = zy = zxx = zxy = zyy = 0;
        for (int a = 0; a < 4; a++)
          for (int b = 0; b < 4; b++) {
            const int dx = a & 1, dy = b & 1;
            const double* src = (a >> 1) ? ((b >> 1) ? dzxy.data() : dzx.data()) : ((b >> 1) ? dzy.data() : zs);
            const double c = at(src, dx, dy);
            z += c * Bx[a] * By[b];
            zx += c * Dx[a] * By[b];
            zy += c * Bx[a] * Dy[b];
            zxx += c * DDx[a] * By[b];
            zxy += c * Dx[a] * Dy[b];
            zyy += c * Bx[a] * DDy[b];
          }
      } else {  // :363-388
        const double z00 = at(zs, 0, 0), z10 = at(zs, 1, 0), z01 = at(zs, 0, 1), z11 = at(zs, 1, 1);
        const double lo = z00 * (1 - xf) + z10 * xf, hi = z01 * (1 - xf) + z11 * xf;
        z = lo * (1 - yf) + hi * yf;
        zx = ((z10 - z00) * (1 - yf) + (z11 - z01) * yf) / hx;
        zy = (hi - lo) / hy;
        zxx = zyy = 0;
        zxy = ((z11 - z01) - (z10 - z00)) / (hx * hy);
      }
      fx[0] = z;
      if (Jp) {
        (*Jp)(0, 0) = zx;
        (*Jp)(0, 1) = zy;
      }
      if (Hp) {
        (*Hp)(0, 0) += l[0] * zxx;
        (*Hp)(0, 1) += l[0] * zxy;
        (*Hp)(1, 0) += l[0] * zxy;
        (*Hp)(1, 1) += l[0] * zyy;
      }
    } break;
    case K_MATPROD: {  // MatrixFunctionProduct (MatrixProduct.h:19-442): C = A B over the children's outputs read as matrices
      const int r = n.ip[0], kk = n.ip[1], c = n.ip[2];
      const bool arow = n.ip[3] != 0, brow = n.ip[4] != 0;
      const int nb = off[1];  // B's outputs follow A's in Y
      for (int j = 0; j < c; j++)
        for (int i = 0; i < r; i++) {
          const int o = j * r + i;  // column-major result
          fx[o] = 0.0;
          for (int q = 0; q < kk; q++) {
            const int ia = arow ? i * kk + q : q * r + i, ib = nb + (brow ? q * c + j : j * kk + q);
            fx[o] += Y[ia] * Y[ib];
            if (Jp) {
              (*Jp)(o, ia) += Y[ib];
              (*Jp)(o, ib) += Y[ia];
            }
            if (Hp) {
              (*Hp)(ia, ib) += l[o];
              (*Hp)(ib, ia) += l[o];
            }
          }
        }
    } break;
    case K_MATINV: {  // MatrixInverse (MatrixInverse.h:6-188): d(A^-1) = -A^-1 dA A^-1, second derivative by the product rule
      const int m = n.ip[0];
      const bool rowm = n.ip[1] != 0;
      auto at = [&](int i, int j) { return rowm ? i * m + j : j * m + i; };
      Mat A(m, 2 * m);
      for (int i = 0; i < m; i++) {
        for (int j = 0; j < m; j++) A(i, j) = Y[at(i, j)];
        A(i, m + i) = 1.0;
      }
      for (int col = 0; col < m; col++) {  // Gauss-Jordan, partial pivoting
        int piv = col;
        for (int i = col + 1; i < m; i++)
          if (std::fabs(A(i, col)) > std::fabs(A(piv, col))) piv = i;
        for (int q = 0; q < 2 * m; q++) std::swap(A(col, q), A(piv, q));
        const double ip = 1.0 / A(col, col);
        for (int q = 0; q < 2 * m; q++) A(col, q) *= ip;
        for (int i = 0; i < m; i++) {
          if (i == col) continue;
          const double f = A(i, col);
          if (f != 0.0)
            for (int q = 0; q < 2 * m; q++) A(i, q) -= f * A(col, q);
        }
      }
      auto inv = [&](int i, int j) { return A(i, m + j); };
      for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) fx[at(i, j)] = inv(i, j);
      if (Jp || Hp)
        for (int a = 0; a < m; a++)
          for (int b = 0; b < m; b++)      // input element A(a, b)
            for (int p = 0; p < m; p++)
              for (int q = 0; q < m; q++) {  // output element Ainv(p, q)
                if (Jp) (*Jp)(at(p, q), at(a, b)) = -inv(p, a) * inv(b, q);
                if (Hp)
                  for (int c2 = 0; c2 < m; c2++)
                    for (int d2 = 0; d2 < m; d2++)
                      (*Hp)(at(a, b), at(c2, d2)) += l[at(p, q)] * (inv(p, a) * inv(b, c2) * inv(d2, q) + inv(p, c2) * inv(d2, a) * inv(b, q));
              }
    } break;
    case K_INTERPND: {  // InterpTable3D / 4D (InterpTable3D.h:122-560, InterpTable4D.h): Hermite patch over the cell, tensor form
      const int D = n.ip[0];
      const bool cubic = n.ip[1] == 0;
      int nn[4], e[4];
      size_t stride[4], cells = 1;
      const double* ax[4];
      const double* p = n.d.data();
      for (int d = 0; d < D; d++) {
        nn[d] = n.ip[2 + d];
        ax[d] = p;
        p += nn[d];
        cells *= (size_t) nn[d];
      }
      for (int d = D - 1, sacc = 1; d >= 0; d--) {
        stride[d] = (size_t) sacc;
        sacc *= nn[d];
      }
      const double* vals = p;
      // mixed derivative arrays recomputed here by successive stencils (calc_derivs), lowest set axis last
      std::call_once(n.aux_once, [&]() {
        n.aux.assign((size_t) (cubic ? 1 << D : 1), Vec());
        n.aux[0].assign(vals, vals + cells);
        if (cubic)
          for (int b = 1; b < (1 << D); b++) {
            int d = 0;
            while (!((b >> d) & 1)) d++;  // differentiate the array without b's lowest bit along that axis
            const Vec& src = n.aux[(size_t) (b & (b - 1))];
            n.aux[(size_t) b].assign(cells, 0.0);
            for (size_t base = 0; base < cells; base++) {
              if ((base / stride[d]) % (size_t) nn[d] != 0) continue;  // first element of each line along axis d
              stencil_axis(ax[d], nn[d], &src[base], &n.aux[(size_t) b][base], stride[d]);
            }
          }
      });
      const std::vector<Vec>& arr = n.aux;
      double B[4][4], D1[4][4], D2[4][4];
      size_t cell = 0;
      for (int d = 0; d < D; d++) {
        e[d] = interval_of(ax[d], nn[d], n.ip[2 + D + d] != 0, Y[d]);
        const double hh = ax[d][e[d] + 1] - ax[d][e[d]], u = (Y[d] - ax[d][e[d]]) / hh;
        cell += (size_t) e[d] * stride[d];
        if (cubic) {
          const double u2 = u * u, u3 = u2 * u;
          B[d][0] = 2 * u3 - 3 * u2 + 1, B[d][1] = -2 * u3 + 3 * u2, B[d][2] = (u3 - 2 * u2 + u) * hh, B[d][3] = (u3 - u2) * hh;
          D1[d][0] = (6 * u2 - 6 * u) / hh, D1[d][1] = -D1[d][0], D1[d][2] = 3 * u2 - 4 * u + 1, D1[d][3] = 3 * u2 - 2 * u;
          D2[d][0] = (12 * u - 6) / (hh * hh), D2[d][1] = -D2[d][0], D2[d][2] = (6 * u - 4) / hh, D2[d][3] = (6 * u - 2) / hh;
        } else {
          B[d][0] = 1 - u, B[d][1] = u, B[d][2] = B[d][3] = 0;
          D1[d][0] = -1 / hh, D1[d][1] = 1 / hh, D1[d][2] = D1[d][3] = 0;
          D2[d][0] = D2[d][1] = D2[d][2] = D2[d][3] = 0;
        }
      }
      double f = 0, g[4] = {0, 0, 0, 0}, H2[4][4] = {{0}};
      const int per = cubic ? 4 : 2;
      int w[4] = {0, 0, 0, 0};
      for (;;) {  // every combination of (slope flag, corner) per axis: w[d] = flag * 2 + corner
        int fb = 0;
        size_t o = cell;
        for (int d = 0; d < D; d++) {
          fb |= (w[d] >> 1) << d;
          o += (size_t) (w[d] & 1) * stride[d];
        }
        const double c = arr[(size_t) fb][o];
        double prod = c;
        for (int d = 0; d < D; d++) prod *= B[d][w[d]];
        f += prod;
        for (int a = 0; a < D; a++) {
          double ga = c;
          for (int d = 0; d < D; d++) ga *= (d == a ? D1[d][w[d]] : B[d][w[d]]);
          g[a] += ga;
          for (int b = a; b < D; b++) {
            double hab = c;
            for (int d = 0; d < D; d++) {
              if (a == b)
                hab *= (d == a ? D2[d][w[d]] : B[d][w[d]]);
              else
                hab *= (d == a || d == b ? D1[d][w[d]] : B[d][w[d]]);
            }
            H2[a][b] += hab;
          }
        }
        int d = 0;
        while (d < D && ++w[d] == per) w[d++] = 0;
        if (d == D) break;
      }
      fx[0] = f;
      for (int a = 0; a < D; a++) {
        if (Jp) (*Jp)(0, a) = g[a];
        if (Hp)
          for (int b = a; b < D; b++) {
            (*Hp)(a, b) += l[0] * H2[a][b];
            if (b != a) (*Hp)(b, a) += l[0] * H2[a][b];
          }
      }
    } break;
    case K_CWISEPROD: {  // CwiseProduct.h
      const int m = off[1];
      for (int i = 0; i < m; i++) {
        fx[i] = Y[i] * Y[m + i];
        if (Jp) {
          (*Jp)(i, i) = Y[m + i];
          (*Jp)(i, m + i) = Y[i];
        }
        if (Hp) {
          (*Hp)(i, m + i) += l[i];
          (*Hp)(m + i, i) += l[i];
        }
      }
    } break;
    case K_CWISEQUOT: {
      const int m = off[1];
      for (int i = 0; i < m; i++) {
        const double a = Y[i], b = Y[m + i];
        fx[i] = a / b;
        if (Jp) {
          (*Jp)(i, i) = 1 / b;
          (*Jp)(i, m + i) = -a / (b * b);
        }
        if (Hp) {
          (*Hp)(i, m + i) += -l[i] / (b * b);
          (*Hp)(m + i, i) += -l[i] / (b * b);
          (*Hp)(m + i, m + i) += 2 * l[i] * a / (b * b * b);
        }
      }
    } break;
    case K_VSPROD: {  // VectorScalarFunctionProduct.h
      const int m = off[1];
      const double s = Y[m];
      for (int i = 0; i < m; i++) {
        fx[i] = Y[i] * s;
        if (Jp) {
          (*Jp)(i, i) = s;
          (*Jp)(i, m) = Y[i];
        }
        if (Hp) {
          (*Hp)(i, m) += l[i];
          (*Hp)(m, i) += l[i];
        }
      }
    } break;
    case K_VSDIV: {  // VectorScalarFunctionDivision.h
      const int m = off[1];
      const double s = Y[m];
      for (int i = 0; i < m; i++) {
        fx[i] = Y[i] / s;
        if (Jp) {
          (*Jp)(i, i) = 1 / s;
          (*Jp)(i, m) = -Y[i] / (s * s);
        }
        if (Hp) {
          (*Hp)(i, m) += -l[i] / (s * s);
          (*Hp)(m, i) += -l[i] / (s * s);
          (*Hp)(m, m) += 2 * l[i] * Y[i] / (s * s * s);
        }
      }
    } break;
    case K_CWISESUM:  // CwiseSum.h
      fx[0] = 0;
      for (int i = 0; i < ny; i++) {
        fx[0] += Y[i];
        if (Jp) (*Jp)(0, i) = 1.0;
      }
      break;
    case K_ATAN2: {  // ArcTan2.h:28-116   f = atan2(y, x)
      const double y = Y[0], x = Y[1], r2 = x * x + y * y;
      fx[0] = std::atan2(y, x);
      if (Jp) {
        (*Jp)(0, 0) = x / r2;
        (*Jp)(0, 1) = -y / r2;
      }
      if (Hp) {
        const double r4 = r2 * r2;
        (*Hp)(0, 0) += l[0] * (-2 * x * y / r4);
        (*Hp)(1, 1) += l[0] * (2 * x * y / r4);
        (*Hp)(0, 1) += l[0] * ((y * y - x * x) / r4);
        (*Hp)(1, 0) += l[0] * ((y * y - x * x) / r4);
      }
    } break;
    case K_COMPARE: {  // Comparative.h: 0/1 with zero derivative
      const double a = Y[0], b = Y[1];
      bool r = false;
      switch (n.ip[0]) {
        case 0: r = a < b; break;
        case 1: r = a <= b; break;
        case 2: r = a > b; break;
        default: r = a >= b; break;
      }
      fx[0] = r ? 1.0 : 0.0;
    } break;
    case K_BOOL: fx[0] = (n.ip[0] == 0) ? ((Y[0] != 0 && Y[1] != 0) ? 1.0 : 0.0) : ((Y[0] != 0 || Y[1] != 0) ? 1.0 : 0.0); break;
    default: throw std::invalid_argument("oracle: outer() called for kind " + std::to_string(n.kind));
  }
}

void Eval::run(const Node& n, const double* x, double* fx, Mat* J, const double* l, Mat* H) const {
  const int ir = n.ir, orr = n.orows;
  if (J) J->reset(orr, ir);
  switch (n.kind) {
    case K_ARGS:  // Arguments.h: identity
      for (int i = 0; i < ir; i++) {
        fx[i] = x[i];
        if (J) (*J)(i, i) = 1.0;
      }
      return;
    case K_CONSTANT:  // Constant.h
      for (int i = 0; i < orr; i++) fx[i] = n.d[i];
      return;
    case K_APPCONST:
      for (int i = 0; i < orr; i++) fx[i] = appc[n.ip[0] + i];
      return;
    case K_NESTED: {  // NestedFunction.h:140-270, ch[0] outer, ch[1] inner
      const Node& o = *n.ch[0];
      const Node& in = *n.ch[1];
      Vec y(in.orows);
      Mat Ji, Jo, Ho;
      const bool deriv = J || H;
      run(in, x, y.data(), deriv ? &Ji : nullptr, nullptr, nullptr);
      if (H) Ho.reset(in.orows, in.orows);
      run(o, y.data(), fx, deriv ? &Jo : nullptr, l, H ? &Ho : nullptr);
      if (J) gemm_acc(*J, Jo, Ji);
      if (H) {
        congruence_acc(*H, Ji, Ho);
        Vec li = atv(Jo, l);
        Mat dummyJ;
        Vec tmp(in.orows);
        run(in, x, tmp.data(), nullptr, li.data(), H);
      }
      return;
    }
    case K_IOSCALED: {  // IOScaled.h:8-158: out .* f(in .* x)
      const Node& f = *n.ch[0];
      Vec xs(ir), y(orr), ls(orr);
      for (int i = 0; i < ir; i++) xs[i] = x[i] * n.d[i];
      Mat Jf, Hf;
      if (H) {
        Hf.reset(ir, ir);
        for (int i = 0; i < orr; i++) ls[i] = l[i] * n.d[ir + i];
      }
      run(f, xs.data(), y.data(), J ? &Jf : nullptr, H ? ls.data() : nullptr, H ? &Hf : nullptr);
      for (int i = 0; i < orr; i++) {
        fx[i] = y[i] * n.d[ir + i];
        if (J)
          for (int j = 0; j < ir; j++) (*J)(i, j) = n.d[ir + i] * Jf(i, j) * n.d[j];
      }
      if (H)
        for (int i = 0; i < ir; i++)
          for (int j = 0; j < ir; j++) (*H)(i, j) += n.d[i] * Hf(i, j) * n.d[j];
      return;
    }
    case K_IFELSE: {  // Conditional.h:19-235: only the taken branch is evaluated / differentiated
      double c;
      run(*n.ch[0], x, &c, nullptr, nullptr, nullptr);
      run(c != 0.0 ? *n.ch[1] : *n.ch[2], x, fx, J, l, H);
      return;
    }
    default: break;
  }
  // generic: phi(stack of children outputs)
  const size_t nc = n.ch.size();
  std::vector<int> off(nc + 1, 0);
  for (size_t k = 0; k < nc; k++) off[k + 1] = off[k] + n.ch[k]->orows;
  const int ny = off[nc];
  Vec Y(ny);
  std::vector<Mat> Jc(nc);
  const bool deriv = J || H;
  for (size_t k = 0; k < nc; k++) run(*n.ch[k], x, &Y[off[k]], deriv ? &Jc[k] : nullptr, nullptr, nullptr);
  Mat Jp, Hp;
  if (deriv) Jp.reset(orr, ny);
  if (H) Hp.reset(ny, ny);
  outer(n, Y, off, fx, deriv ? &Jp : nullptr, l, H ? &Hp : nullptr);
  if (!deriv) return;
  Mat Js(ny, ir);
  for (size_t k = 0; k < nc; k++)
    for (int i = 0; i < n.ch[k]->orows; i++)
      for (int j = 0; j < ir; j++) Js(off[k] + i, j) = Jc[k](i, j);
  if (J) gemm_acc(*J, Jp, Js);
  if (H) {
    congruence_acc(*H, Js, Hp);
    Vec ls = atv(Jp, l);
    Vec tmp;
    for (size_t k = 0; k < nc; k++) {
      tmp.assign(n.ch[k]->orows, 0.0);
      run(*n.ch[k], x, tmp.data(), nullptr, &ls[off[k]], H);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// LGL coefficient tables: decimal literals of src/OptimalControl/LGLCoeffs.h:14-393
// ------------------------------------------------------------------------------------------------
struct Coeffs {
  int CS;
  double card[4], inter[3], XI[3][4], DXI[3][4], UP[3][4], XD[3][4], DXD[3][4], IDXD[3], RW[4], UZ[2][4], UO[2][4];
};
static Coeffs coeffs(int CS) {
  Coeffs c;
  std::memset(&c, 0, sizeof(c));
  c.CS = CS;
  if (CS == 2) {  // LGLCoeffs<2>, LGLCoeffs.h:14-56
    c.card[1] = 1.0;
    c.inter[0] = 0.5;
    c.XI[0][0] = c.XI[0][1] = 0.5;
    c.UP[0][0] = c.UP[0][1] = 0.5;
    c.DXI[0][0] = 1.0 / 8.0;
    c.DXI[0][1] = -1.0 / 8.0;
    c.XD[0][0] = 1.0;
    c.XD[0][1] = -1.0;
    c.DXD[0][0] = c.DXD[0][1] = 1.0 / 6.0;
    c.IDXD[0] = 4.0 / 6.0;
    c.RW[0] = c.RW[1] = 0.5;
  } else if (CS == 3) {  // LGLCoeffs<3>, LGLCoeffs.h:61-162
    const double s = 4.58257569495584;
    c.card[1] = 0.5;
    c.card[2] = 1.0;
    c.inter[0] = 0.172673164646011;
    c.inter[1] = 0.827326835353989;
    const double q = 1.0 / 686.0, w = 1.0 / 360.0;
    const double xi[2][3] = {{(39 * s + 231) * q, 224 * q, (-39 * s + 231) * q}, {(-39 * s + 231) * q, 224 * q, (39 * s + 231) * q}};
    const double dxi[2][3] = {{(3 * s + 21) * q, -16 * s * q, (3 * s - 21) * q}, {(-3 * s + 21) * q, 16 * s * q, (-3 * s - 21) * q}};
    const double xd[2][3] = {{(32 * s + 180) * w, -64 * s * w, (32 * s - 180) * w}, {(-32 * s + 180) * w, 64 * s * w, (-32 * s - 180) * w}};
    const double dxd[2][3] = {{(9 + s) * w, 64 * w, (9 - s) * w}, {(9 - s) * w, 64 * w, (9 + s) * w}};
    const double up[2][3] = {{0.541612549639704, 0.571428571428571, -0.113041121068274},
                             {-0.113041121068274, 0.571428571428571, 0.541612549639704}};
    for (int i = 0; i < 2; i++)
      for (int j = 0; j < 3; j++) {
        c.XI[i][j] = xi[i][j];
        c.DXI[i][j] = dxi[i][j];
        c.XD[i][j] = xd[i][j];
        c.DXD[i][j] = dxd[i][j];
        c.UP[i][j] = up[i][j];
      }
    c.IDXD[0] = c.IDXD[1] = 98 * w;
    c.RW[0] = 1.0 / 6.0;
    c.RW[1] = 2.0 / 3.0;
    c.RW[2] = 1.0 / 6.0;
    const double uz[3] = {-3, 4, -1}, uo[3] = {1, -4, 3};
    for (int j = 0; j < 3; j++) {
      c.UZ[0][j] = uz[j];
      c.UO[0][j] = uo[j];
    }
  } else if (CS == 4) {  // LGLCoeffs<4>, LGLCoeffs.h:168-393
    const double card[4] = {0.0, 2.65575603264643e-1, 7.34424396735357e-1, 1.0};
    const double inter[3] = {8.48880518607166e-2, 0.5, 9.15111948139283e-1};
    const double xi[3][4] = {{6.18612232711785e-1, 3.34253095933642e-1, 1.52679626438851e-2, 3.18667087106879e-2},
                             {1.41445282326366e-1, 3.58554717673634e-1, 3.58554717673634e-1, 1.41445282326366e-1},
                             {3.18667087106879e-2, 1.52679626438851e-2, 3.34253095933642e-1, 6.18612232711785e-1}};
    const double dxi[3][4] = {{2.57387738427162e-2, -5.50098654524528e-2, -1.53026046503702e-2, -2.38759243962924e-3},
                              {9.92317607754556e-3, 9.62835932121973e-2, -9.62835932121973e-2, -9.92317607754556e-3},
                              {2.38759243962924e-3, 1.53026046503702e-2, 5.50098654524528e-2, -2.57387738427162e-2}};
    const double up[3][4] = {{0.550643660407289, 0.551767574740443, -0.153490305524281, 0.0510790703765507},
                             {-0.140877081724073, 0.640877081724073, 0.640877081724073, -0.140877081724073},
                             {0.0510790703765507, -0.153490305524281, 0.551767574740443, 0.550643660407289}};
    const double xd[3][4] = {{8.84260109348311e-1, -8.23622559094327e-1, -2.35465327970606e-2, -3.70910174569208e-2},
                             {7.86488731947674e-2, 8.00076026297266e-1, -8.00076026297266e-1, -7.86488731947674e-2},
                             {3.70910174569208e-2, 2.35465327970606e-2, 8.23622559094327e-1, -8.84260109348311e-1}};
    const double dxd[3][4] = {{1.62213410652341e-2, 9.71662045547156e-2, 1.85682012187242e-2, 2.74945307600086e-3},
                              {4.83872966828888e-3, 1.00138284831491e-1, 1.00138284831491e-1, 4.83872966828888e-3},
                              {2.74945307600086e-3, 1.85682012187242e-2, 9.71662045547156e-2, 1.62213410652341e-2}};
    for (int j = 0; j < 4; j++) c.card[j] = card[j];
    for (int i = 0; i < 3; i++) {
      c.inter[i] = inter[i];
      for (int j = 0; j < 4; j++) {
        c.XI[i][j] = xi[i][j];
        c.DXI[i][j] = dxi[i][j];
        c.UP[i][j] = up[i][j];
        c.XD[i][j] = xd[i][j];
        c.DXD[i][j] = dxd[i][j];
      }
    }
    c.IDXD[0] = 1.38413023680783e-1;
    c.IDXD[1] = 2.43809523809524e-1;
    c.IDXD[2] = 1.38413023680783e-1;
    // UPoly_Weights of LGLCoeffs<4> (LGLCoeffs.h:300-364): cubic Lagrange basis coefficients on the cardinal nodes
    const double a3[4] = {-5.12701665379258, 10.9353308042859, -10.9353308042859, 5.12701665379258};
    const double a2[4] = {10.2540333075852, -18.9665045333251, 13.8394878795326, -5.12701665379258};
    const double a1[4] = {-6.12701665379258, 8.03117372903925, -2.90415707524666, 1.0};
    const double a0[4] = {1.0, 0.0, 0.0, 0.0};
    for (int j = 0; j < 4; j++) {
      c.RW[j] = a3[j] / 4.0 + a2[j] / 3.0 + a1[j] / 2.0 + a0[j];
      c.UZ[0][j] = a1[j];
      c.UZ[1][j] = a2[j] * 2.0;
      c.UO[0][j] = a3[j] * 3.0 + a2[j] * 2.0 + a1[j];
      c.UO[1][j] = a3[j] * 6.0 + a2[j] * 2.0;
    }
  } else {
    throw std::invalid_argument("oracle: CS must be 2, 3 or 4");
  }
  return c;
}

// ------------------------------------------------------------------------------------------------
// per-application functions
// ------------------------------------------------------------------------------------------------
enum FnKind { F_EXPR, F_LGL_DEFECT, F_TRAP_DEFECT, F_LGL_MESH, F_SINGLE_MESH, F_SPLINE, F_INTEGRAL };

struct Function {
  int fkind = F_EXPR;
  int ir = 0, orows = 0;
  int CS = 0, XV = 0, UV = 0, PV = 0, order = 0;
  std::unique_ptr<Node> expr;  // ODE, integrand or the function itself
  Coeffs cf;

  // fx (orows), jx (orows x ir), hx (ir x ir) are assigned.  want: 0 values, 1 +jacobian, 2 +hessian
  void compute(const double* x, const double* l, const double* appc, double* fx, Mat& jx, Mat& hx, int want) const;

 private:
  void ode(const double* X, const double* l, double* f, Mat* J, Vec* ag, Mat* H) const {
    Eval ev;
    const int no = expr->ir;
    if (H) H->reset(no, no);
    ev.run(*expr, X, f, J, l, H);
    if (ag) *ag = atv(*J, l);
  }
  void lgl_defect(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const;
  void trap_defect(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const;
  void integral(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const;
  void spline(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const;
  void lgl_mesh(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const;
};

// LGLDefects<DODE,CS>::compute_jacobian_adjointgradient_adjointhessian_impl, LGLDefects.h:289-551.
// Variable names follow the reference (C_XS, C_DXS, C_JDXS, C_AVS, I_*, DI_DCS, HTpar).
void Function::lgl_defect(const double* x, const double* adjvars, double* fx, Mat& jx, Mat& adjhess, int want) const {
  const int n = XV + 1 + UV, no = n + PV, T = XV, IR = ir;
  const int Card = CS, Int = CS - 1;
  std::vector<Vec> C_XS(Card, Vec(no)), C_DXS(Card, Vec(XV)), I_XS(Int, Vec(no, 0.0)), I_DXS(Int, Vec(XV)), I_AVS(Int, Vec(XV)),
      C_AVS(Card, Vec(XV, 0.0)), I_AGXS(Int);
  std::vector<Mat> C_JDXS(Card), I_JDXS(Int), I_HDXS(Int);
  for (int i = 0; i < Card; i++) {  // :322-334
    for (int k = 0; k < n; k++) C_XS[i][k] = x[i * n + k];
    for (int k = 0; k < PV; k++) C_XS[i][n + k] = x[Card * n + k];
    ode(C_XS[i].data(), nullptr, C_DXS[i].data(), nullptr, nullptr, nullptr);
  }
  const double h = C_XS[Card - 1][T] - C_XS[0][T];  // :336
  for (int i = 0; i < Int; i++) {                     // :338-375
    I_XS[i][T] = C_XS[0][T] + h * cf.inter[i];
    for (int k = 0; k < PV; k++) I_XS[i][n + k] = x[Card * n + k];
    for (int k = 0; k < XV; k++) I_AVS[i][k] = adjvars ? adjvars[i * XV + k] : 0.0;
    for (int j = 0; j < Card; j++) {
      for (int k = 0; k < XV; k++) I_XS[i][k] += cf.XI[i][j] * C_XS[j][k] + (cf.DXI[i][j] * h) * C_DXS[j][k];
      for (int k = 0; k < UV; k++) I_XS[i][XV + 1 + k] += cf.UP[i][j] * C_XS[j][XV + 1 + k];
    }
    if (want == 0) {
      ode(I_XS[i].data(), nullptr, I_DXS[i].data(), nullptr, nullptr, nullptr);
    } else {
      ode(I_XS[i].data(), I_AVS[i].data(), I_DXS[i].data(), &I_JDXS[i], &I_AGXS[i], want >= 2 ? &I_HDXS[i] : nullptr);
      for (int j = 0; j < Card; j++) {
        const double scale = cf.IDXD[i] * cf.DXI[i][j];
        for (int k = 0; k < XV; k++) {
          C_AVS[j][k] += I_AGXS[i][k] * (scale * h * h);
          C_AVS[j][k] += I_AVS[i][k] * (cf.DXD[i][j] * h);
        }
      }
    }
  }
  for (int k = 0; k < orows; k++) fx[k] = 0.0;
  if (want == 0) {  // compute_impl, :57-122
    for (int i = 0; i < Int; i++) {
      for (int j = 0; j < Card; j++)
        for (int k = 0; k < XV; k++) fx[i * XV + k] += cf.XD[i][j] * C_XS[j][k] + (cf.DXD[i][j] * h) * C_DXS[j][k];
      for (int k = 0; k < XV; k++) fx[i * XV + k] += h * cf.IDXD[i] * I_DXS[i][k];
    }
    return;
  }
  jx.reset(orows, IR);
  adjhess.reset(IR, IR);
  Vec HTpar(IR, 0.0);
  for (int j = 0; j < Card; j++) {  // :377-412
    Vec C_AGXS;
    Mat C_HDXS;
    Vec f(XV);
    ode(C_XS[j].data(), C_AVS[j].data(), f.data(), &C_JDXS[j], &C_AGXS, want >= 2 ? &C_HDXS : nullptr);
    if (want >= 2) {
      for (int a = 0; a < n; a++) {
        for (int b = 0; b < n; b++) adjhess(j * n + a, j * n + b) += C_HDXS(a, b);
        for (int b = 0; b < PV; b++) {
          adjhess(j * n + a, Card * n + b) += C_HDXS(a, n + b);
          adjhess(Card * n + b, j * n + a) += C_HDXS(n + b, a);
        }
      }
      for (int a = 0; a < PV; a++)
        for (int b = 0; b < PV; b++) adjhess(Card * n + a, Card * n + b) += C_HDXS(n + a, n + b);
    }
    for (int a = 0; a < n; a++) HTpar[j * n + a] += C_AGXS[a] * (1.0 / h);
    for (int a = 0; a < PV; a++) HTpar[Card * n + a] += C_AGXS[n + a] * (1.0 / h);
  }
  const int TL = T + n * (Card - 1);
  for (int i = 0; i < Int; i++) {  // :414-513
    Mat DI_DCS(no, IR);
    DI_DCS(T, T) = 1.0 - cf.inter[i];
    DI_DCS(T, TL) = cf.inter[i];
    for (int k = 0; k < PV; k++) DI_DCS(n + k, Card * n + k) = 1.0;
    for (int j = 0; j < Card; j++) {
      for (int k = 0; k < XV; k++) DI_DCS(k, j * n + k) = cf.XI[i][j];
      const double vh = cf.DXI[i][j] * h, wh = cf.DXD[i][j] * h;
      for (int k = 0; k < XV; k++) {
        for (int b = 0; b < n; b++) DI_DCS(k, j * n + b) += vh * C_JDXS[j](k, b);
        for (int b = 0; b < PV; b++) DI_DCS(k, Card * n + b) += vh * C_JDXS[j](k, n + b);
        DI_DCS(k, T) -= cf.DXI[i][j] * C_DXS[j][k];
        DI_DCS(k, TL) += cf.DXI[i][j] * C_DXS[j][k];
      }
      for (int k = 0; k < UV; k++) DI_DCS(XV + 1 + k, j * n + XV + 1 + k) = cf.UP[i][j];
      for (int k = 0; k < XV; k++) {
        fx[i * XV + k] += cf.XD[i][j] * C_XS[j][k] + wh * C_DXS[j][k];
        jx(i * XV + k, j * n + k) = cf.XD[i][j];
      }
      for (int k = 0; k < XV; k++) {
        for (int b = 0; b < n; b++) jx(i * XV + k, j * n + b) += wh * C_JDXS[j](k, b);
        for (int b = 0; b < PV; b++) jx(i * XV + k, Card * n + b) += wh * C_JDXS[j](k, n + b);
        jx(i * XV + k, T) -= cf.DXD[i][j] * C_DXS[j][k];
        jx(i * XV + k, TL) += cf.DXD[i][j] * C_DXS[j][k];
      }
    }
    const double hw = h * cf.IDXD[i];
    for (int k = 0; k < XV; k++) fx[i * XV + k] += hw * I_DXS[i][k];
    {  // jx.middleRows += (hw * I_JDXS) * DI_DCS
      Mat blk(XV, IR);
      gemm_acc(blk, I_JDXS[i], DI_DCS, hw);
      for (int k = 0; k < XV; k++)
        for (int b = 0; b < IR; b++) jx(i * XV + k, b) += blk(k, b);
    }
    for (int k = 0; k < XV; k++) {
      jx(i * XV + k, T) -= cf.IDXD[i] * I_DXS[i][k];
      jx(i * XV + k, TL) += cf.IDXD[i] * I_DXS[i][k];
    }
    if (want >= 2) {
      congruence_acc(adjhess, DI_DCS, I_HDXS[i], hw);
      for (int b = 0; b < IR; b++) {
        double s = 0;
        for (int a = 0; a < no; a++) s += I_AGXS[i][a] * cf.IDXD[i] * DI_DCS(a, b);
        HTpar[b] += s;
      }
    }
  }
  if (want >= 2) {  // :507-511
    for (int a = 0; a < IR; a++) {
      adjhess(a, T) -= HTpar[a];
      adjhess(a, TL) += HTpar[a];
    }
    for (int a = 0; a < IR; a++) {
      adjhess(T, a) -= HTpar[a];
      adjhess(TL, a) += HTpar[a];
    }
  }
}

// TrapezoidalDefects, TrapezoidalDefects.h:146-435: d = -[(x1 - x0) - (h/2)(f0 + f1)] (note the final sign flip :431-434)
void Function::trap_defect(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const {
  const int n = XV + 1 + UV, no = n + PV, T = XV, IR = ir;
  Vec X0(no), X1(no), F0(XV), F1(XV);
  for (int k = 0; k < n; k++) {
    X0[k] = x[k];
    X1[k] = x[n + k];
  }
  for (int k = 0; k < PV; k++) X0[n + k] = X1[n + k] = x[2 * n + k];
  const double h = X1[T] - X0[T];
  Mat J0, J1, H0, H1;
  Vec G0, G1;
  Vec lz(XV, 0.0);
  const double* lam = l ? l : lz.data();
  if (want == 0) {
    ode(X0.data(), nullptr, F0.data(), nullptr, nullptr, nullptr);
    ode(X1.data(), nullptr, F1.data(), nullptr, nullptr, nullptr);
  } else {
    ode(X0.data(), lam, F0.data(), &J0, &G0, want >= 2 ? &H0 : nullptr);
    ode(X1.data(), lam, F1.data(), &J1, &G1, want >= 2 ? &H1 : nullptr);
  }
  for (int k = 0; k < XV; k++) fx[k] = -((X1[k] - X0[k]) - 0.5 * h * (F0[k] + F1[k]));
  if (want == 0) return;
  jx.reset(orows, IR);
  hx.reset(IR, IR);
  // positive form first: d+ = (x1 - x0) - (h/2)(f0 + f1), flipped at the end
  for (int k = 0; k < XV; k++) {
    jx(k, k) += -1.0;
    jx(k, n + k) += 1.0;
    for (int b = 0; b < n; b++) {
      jx(k, b) += -0.5 * h * J0(k, b);
      jx(k, n + b) += -0.5 * h * J1(k, b);
    }
    for (int b = 0; b < PV; b++) jx(k, 2 * n + b) += -0.5 * h * (J0(k, n + b) + J1(k, n + b));
    jx(k, T) += 0.5 * (F0[k] + F1[k]);
    jx(k, n + T) += -0.5 * (F0[k] + F1[k]);
  }
  if (want >= 2) {
    auto map0 = [&](int a) { return a < n ? a : 2 * n + (a - n); };
    auto map1 = [&](int a) { return a < n ? n + a : 2 * n + (a - n); };
    for (int a = 0; a < no; a++)
      for (int b = 0; b < no; b++) {
        hx(map0(a), map0(b)) += -0.5 * h * H0(a, b);
        hx(map1(a), map1(b)) += -0.5 * h * H1(a, b);
      }
    // d/dh coupling: -(1/2) (G0 + G1) against dh/dt0 = -1, dh/dt1 = +1
    Vec HT(IR, 0.0);
    for (int a = 0; a < no; a++) {
      HT[map0(a)] += -0.5 * G0[a];
      HT[map1(a)] += -0.5 * G1[a];
    }
    for (int a = 0; a < IR; a++) {
      hx(a, T) -= HT[a];
      hx(a, n + T) += HT[a];
      hx(T, a) -= HT[a];
      hx(n + T, a) += HT[a];
    }
  }
  for (double& v: jx.a) v = -v;
  for (double& v: hx.a) v = -v;
}

// LGLIntegral = (sum_i W_i g([x_i ; p])) * h, LGLIntegrals.h:9-73 (expression template: product rule on sum * h)
void Function::integral(const double* x, const double* l, double* fx, Mat& jx, Mat& hx, int want) const {
  const int xv = XV, pv = PV, xtv = xv + 1, IR = ir;
  const int t0 = xv, t1 = (CS - 1) * xtv + xv;
  const double h = x[t1] - x[t0];
  double S = 0;
  Vec gS(IR, 0.0);
  if (want >= 1) {
    jx.reset(1, IR);
    hx.reset(IR, IR);
  }
  Eval ev;
  for (int i = 0; i < CS; i++) {
    Vec xi(xv + pv);
    for (int k = 0; k < xv; k++) xi[k] = x[i * xtv + k];
    for (int k = 0; k < pv; k++) xi[xv + k] = x[CS * xtv + k];
    double g;
    Mat Jg, Hg;
    const double lw = (l ? l[0] : 0.0) * h * cf.RW[i];
    if (want >= 2) Hg.reset(xv + pv, xv + pv);
    ev.run(*expr, xi.data(), &g, want >= 1 ? &Jg : nullptr, &lw, want >= 2 ? &Hg : nullptr);
    S += cf.RW[i] * g;
    if (want >= 1) {
      auto map = [&](int a) { return a < xv ? i * xtv + a : CS * xtv + (a - xv); };
      for (int a = 0; a < xv + pv; a++) {
        gS[map(a)] += cf.RW[i] * Jg(0, a);
        if (want >= 2)
          for (int b = 0; b < xv + pv; b++) hx(map(a), map(b)) += Hg(a, b);
      }
    }
  }
  fx[0] = S * h;
  if (want == 0) return;
  for (int a = 0; a < IR; a++) jx(0, a) = gS[a] * h;
  jx(0, t0) -= S;
  jx(0, t1) += S;
  if (want >= 2) {
    const double lam = l[0];
    for (int a = 0; a < IR; a++) {
      hx(a, t0) -= lam * gS[a];
      hx(t0, a) -= lam * gS[a];
      hx(a, t1) += lam * gS[a];
      hx(t1, a) += lam * gS[a];
    }
  }
}

// LGLControlSpline<CSC,-1,Order>, LGLControlSplines.h:64-312
void Function::spline(const double* x, const double* adjvars, double* fx, Mat& jx, Mat& adjhess, int want) const {
  const int CSC = CS, U = UV, tU = U + 1, IR = ir;
  auto t = [&](int i) { return x[i * tU]; };
  auto u = [&](int i, int k) { return x[i * tU + 1 + k]; };
  const double h0 = t(CSC - 1) - t(0), h1 = t(2 * CSC - 2) - t(CSC - 1);
  const int c0 = 0, c1 = (CSC - 1) * tU, c2 = (2 * CSC - 2) * tU;
  for (int k = 0; k < orows; k++) fx[k] = 0;
  if (want >= 1) {
    jx.reset(orows, IR);
    adjhess.reset(IR, IR);
  }
  for (int j = 0; j < order; j++) {
    const double h0pow = 1.0 / std::pow(h0, (double) (j + 1)), h1pow = 1.0 / std::pow(h1, (double) (j + 1));
    const double hdt = j + 1, h2dt = j + 2;
    double OTH = 0, ZTH = 0;
    for (int i = 0; i < CSC; i++) {
      const double wo = cf.UO[j][i], wz = cf.UZ[j][i];
      for (int k = 0; k < U; k++) {
        const int row = j * U + k;
        fx[row] += (wo * h0pow) * u(i, k) - (wz * h1pow) * u(i + CSC - 1, k);
        if (want == 0) continue;
        jx(row, c0) += (wo * hdt * h0pow / h0) * u(i, k);
        jx(row, c1) += -(wo * hdt * h0pow / h0) * u(i, k) - (wz * hdt * h1pow / h1) * u(i + CSC - 1, k);
        jx(row, c2) += (wz * hdt * h1pow / h1) * u(i + CSC - 1, k);
        jx(row, i * tU + 1 + k) += wo * h0pow;
        jx(row, (i + CSC - 1) * tU + 1 + k) += -wz * h1pow;
        if (want >= 2) {
          const double lam = adjvars[row];
          OTH += (wo * hdt * h2dt * h0pow / (h0 * h0)) * u(i, k) * lam;
          ZTH += (wz * hdt * h2dt * h1pow / (h1 * h1)) * u(i + CSC - 1, k) * lam;
          const double a = lam * wo * hdt * h0pow / h0, b = lam * wz * hdt * h1pow / h1;
          const int ui = i * tU + 1 + k, vi = (i + CSC - 1) * tU + 1 + k;
          adjhess(c0, ui) += a;
          adjhess(c1, ui) -= a;
          adjhess(ui, c0) += a;
          adjhess(ui, c1) -= a;
          adjhess(c1, vi) -= b;
          adjhess(c2, vi) += b;
          adjhess(vi, c1) -= b;
          adjhess(vi, c2) += b;
        }
      }
    }
    if (want >= 2) {
      adjhess(c0, c0) += OTH;
      adjhess(c1, c1) += OTH;
      adjhess(c0, c1) += -OTH;
      adjhess(c1, c0) += -OTH;
      adjhess(c1, c1) -= ZTH;
      adjhess(c2, c2) -= ZTH;
      adjhess(c2, c1) += ZTH;
      adjhess(c1, c2) += ZTH;
    }
  }
}

// LGLMeshSpacing<CSC>, MeshSpacingConstraints.h:107-193
void Function::lgl_mesh(const double* x, const double* adjvars, double* fx, Mat& jx, Mat& adjhess, int want) const {
  const int CSC = CS;
  const double h = x[CSC - 1] - x[0], h2 = h * h, h3 = h2 * h;
  if (want >= 1) {
    jx.reset(orows, ir);
    adjhess.reset(ir, ir);
  }
  for (int i = 0; i < CSC - 2; i++) {
    const double dx = x[1 + i] - x[0];
    fx[i] = cf.card[i + 1] - dx / h;
    if (want == 0) continue;
    jx(i, i + 1) = -1.0 / h;
    jx(i, 0) = 1.0 / h - dx / h2;
    jx(i, CSC - 1) = dx / h2;
    if (want >= 2) {
      const double a = adjvars[i];
      adjhess(0, i + 1) += -a / h2;
      adjhess(i + 1, 0) += -a / h2;
      adjhess(CSC - 1, i + 1) += a / h2;
      adjhess(i + 1, CSC - 1) += a / h2;
      adjhess(0, 0) += a * 2.0 / h2 - a * 2.0 * dx / h3;
      adjhess(CSC - 1, CSC - 1) += -a * 2.0 * dx / h3;
      adjhess(0, CSC - 1) += a * (2.0 * dx / h3 - 1.0 / h2);
      adjhess(CSC - 1, 0) += a * (2.0 * dx / h3 - 1.0 / h2);
    }
  }
}

void Function::compute(const double* x, const double* l, const double* appc, double* fx, Mat& jx, Mat& hx, int want) const {
  switch (fkind) {
    case F_LGL_DEFECT: lgl_defect(x, l, fx, jx, hx, want); return;
    case F_TRAP_DEFECT: trap_defect(x, l, fx, jx, hx, want); return;
    case F_INTEGRAL: integral(x, l, fx, jx, hx, want); return;
    case F_SPLINE: spline(x, l, fx, jx, hx, want); return;
    case F_LGL_MESH: lgl_mesh(x, l, fx, jx, hx, want); return;
    case F_SINGLE_MESH: {  // SingleMeshSpacing, MeshSpacingConstraints.h:8-103: c*(t2 - t0) - (t1 - t0), linear
      const double c = appc[0];
      fx[0] = c * (x[2] - x[0]) - (x[1] - x[0]);
      if (want >= 1) {
        jx.reset(1, 3);
        hx.reset(3, 3);
        jx(0, 0) = 1.0 - c;
        jx(0, 1) = -1.0;
        jx(0, 2) = c;
      }
      return;
    }
    default: {
      Eval ev;
      ev.appc = appc;
      if (want >= 2) hx.reset(ir, ir);
      ev.run(*expr, x, fx, want >= 1 ? &jx : nullptr, l, want >= 2 ? &hx : nullptr);
      if (want == 1) hx.reset(ir, ir);
      return;
    }
  }
}

static std::unique_ptr<Function> parse_function(const std::string& desc) {
  auto f = std::make_unique<Function>();
  std::istringstream is(desc);
  std::string head;
  is >> head;
  auto rest = [&]() {
    std::string r;
    std::getline(is, r, '\0');
    return r;
  };
  auto parse_expr = [&](const std::string& s) {
    Parser p(s);
    return p.node();
  };
  if (head == "lgl_defect" || head == "trap_defect") {
    is >> f->CS >> f->XV >> f->UV >> f->PV;
    f->expr = parse_expr(rest());
    const int n = f->XV + 1 + f->UV;
    if (head == "lgl_defect") {
      f->fkind = F_LGL_DEFECT;
      f->cf = coeffs(f->CS);
      f->ir = f->CS * n + f->PV;
      f->orows = (f->CS - 1) * f->XV;
    } else {
      f->fkind = F_TRAP_DEFECT;
      f->CS = 2;
      f->ir = 2 * n + f->PV;
      f->orows = f->XV;
    }
    if (f->expr->ir != n + f->PV || f->expr->orows != f->XV) throw std::invalid_argument("oracle: ODE size mismatch");
  } else if (head == "lgl_mesh_spacing") {
    is >> f->CS;
    f->fkind = F_LGL_MESH;
    f->cf = coeffs(f->CS);
    f->ir = f->CS;
    f->orows = f->CS - 2;
  } else if (head == "single_mesh_spacing") {
    f->fkind = F_SINGLE_MESH;
    f->ir = 3;
    f->orows = 1;
  } else if (head == "control_spline") {
    is >> f->CS >> f->UV >> f->order;
    f->fkind = F_SPLINE;
    f->cf = coeffs(f->CS);
    f->ir = (2 * f->CS - 1) * (f->UV + 1);
    f->orows = f->UV * f->order;
  } else if (head == "lgl_integral") {
    is >> f->CS >> f->XV >> f->PV;
    f->expr = parse_expr(rest());
    f->fkind = F_INTEGRAL;
    f->cf = coeffs(f->CS);
    f->ir = f->CS * (f->XV + 1) + f->PV;
    f->orows = 1;
  } else if (head == "expr") {
    f->expr = parse_expr(rest());
    f->fkind = F_EXPR;
    f->ir = f->expr->ir;
    f->orows = f->expr->orows;
  } else {
    throw std::invalid_argument("oracle: unknown function description '" + head + "'");
  }
  return f;
}

// ------------------------------------------------------------------------------------------------
// NLP: function list, sparsity analysis, evaluation
// ------------------------------------------------------------------------------------------------
struct FuncEntry {
  int role = 1;  // 0 objective, 1 equality, 2 inequality
  bool linear = false;
  int nappl = 0, nappconst = 0;
  std::shared_ptr<Function> fn;
  std::vector<int> V, C;
  Vec appconst;
  int triplets() const {
    return (linear ? 0 : fn->ir * (fn->ir + 1) / 2) + (role == 0 ? 0 : fn->ir * fn->orows);
  }
};

struct Problem {
  int nvars = 0, neq = 0, niq = 0, kktdim = 0;
  std::vector<FuncEntry> funcs;
  std::vector<int> order;
  std::vector<long long> trip_base;
  long long n_user = 0, n_solver = 0;
  std::vector<int> indptr, indices, kktloc;
  std::string err;

  void analyze();
  void eval(int threads, int flags, double obj_scale, const double* X, const double* LE, const double* LI,
            const double* solver_coeffs, double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt) const;
};

// getMATSpace + analyzeSparsity (NonLinearProgram.cpp:111-175, :267-344) at Threads = 1: triplets in emission order,
// flipped into the upper triangle, summed into a row-major compressed matrix; KKTLocations by search within the row.
void Problem::analyze() {
  const int PV = nvars, SV = niq, EC = neq, IC = niq;
  kktdim = PV + SV + EC + IC;
  order.clear();
  for (int role = 0; role < 3; role++)
    for (int i = 0; i < (int) funcs.size(); i++)
      if (funcs[i].role == role) order.push_back(i);
  std::vector<std::pair<int, int>> trip;
  trip_base.assign(funcs.size(), 0);
  for (int fi: order) {
    const FuncEntry& f = funcs[fi];
    trip_base[fi] = (long long) trip.size();
    const int conoff = (f.role == 1) ? PV + SV : PV + SV + EC;  // NonLinearProgram.cpp:121-122
    const int ir = f.fn->ir, orr = f.fn->orows;
    for (int a = 0; a < f.nappl; a++) {  // DenseFunctionBase::getKKTSpace, DenseFunctionBase.h:1097-1129
      const int* v = &f.V[(size_t) a * ir];
      const int* c = f.role ? &f.C[(size_t) a * orr] : nullptr;
      for (int i = 0; i < ir; i++) {
        if (!f.linear)
          for (int j = i; j < ir; j++) trip.emplace_back(v[j], v[i]);
        if (f.role)
          for (int j = 0; j < orr; j++) trip.emplace_back(c[j] + conoff, v[i]);
      }
    }
  }
  n_user = (long long) trip.size();
  // solver-owned entries, NonLinearProgram.cpp:198-265
  for (int i = 0; i < SV; i++) trip.emplace_back(PV + SV + EC + i, PV + i);
  for (int i = 0; i < PV; i++) trip.emplace_back(i, i);
  for (int i = 0; i < SV; i++) trip.emplace_back(PV + i, PV + i);
  for (int i = 0; i < EC; i++) trip.emplace_back(PV + SV + i, PV + SV + i);
  for (int i = 0; i < IC; i++) trip.emplace_back(PV + SV + EC + i, PV + SV + EC + i);
  n_solver = (long long) trip.size() - n_user;
  for (auto& t: trip)
    if (t.first > t.second) std::swap(t.first, t.second);  // :296-304 flip to upper triangle
  // Row-major compressed pattern = the sorted, de-duplicated triplets.  Done row by row (counting sort on the row, then a
  // sort of each row's few columns) and over several threads: the same result as sorting all triplets at once, which at
  // 200 k segments means 2.7e8 pairs.
  const int nth = std::max(1, std::min<int>(16, (int) std::thread::hardware_concurrency()));
  std::vector<long long> rstart((size_t) kktdim + 1, 0);
  for (const auto& t: trip) rstart[(size_t) t.first + 1]++;
  for (int r = 0; r < kktdim; r++) rstart[(size_t) r + 1] += rstart[r];
  std::vector<int> cols(trip.size());
  {
    std::vector<long long> fill(rstart.begin(), rstart.end() - 1);
    for (const auto& t: trip) cols[(size_t) fill[(size_t) t.first]++] = t.second;
  }
  std::vector<int> rcount((size_t) kktdim, 0);
  auto rows_of = [&](int w) { return std::make_pair((long long) kktdim * w / nth, (long long) kktdim * (w + 1) / nth); };
  {
    std::vector<std::thread> pool;
    for (int w = 0; w < nth; w++)
      pool.emplace_back([&, w]() {
        const auto rr = rows_of(w);
        for (long long r = rr.first; r < rr.second; r++) {
          int* b = cols.data() + rstart[(size_t) r];
          int* e = cols.data() + rstart[(size_t) r + 1];
          std::sort(b, e);
          rcount[(size_t) r] = (int) (std::unique(b, e) - b);
        }
      });
    for (auto& t: pool) t.join();
  }
  indptr.assign((size_t) kktdim + 1, 0);
  for (int r = 0; r < kktdim; r++) indptr[(size_t) r + 1] = indptr[r] + rcount[(size_t) r];
  indices.resize((size_t) indptr[(size_t) kktdim]);
  for (int r = 0; r < kktdim; r++) std::copy(cols.begin() + rstart[(size_t) r], cols.begin() + rstart[(size_t) r] + rcount[(size_t) r], indices.begin() + indptr[r]);
  std::vector<int>().swap(cols);
  kktloc.resize(trip.size());
  {
    std::vector<std::thread> pool;
    for (int w = 0; w < nth; w++)
      pool.emplace_back([&, w]() {  // :318-335: the position of (row, column) within its row
        const size_t k0 = trip.size() * (size_t) w / (size_t) nth, k1 = trip.size() * (size_t) (w + 1) / (size_t) nth;
        for (size_t k = k0; k < k1; k++) {
          const int r = trip[k].first, c = trip[k].second;
          const int* b = indices.data() + indptr[r];
          const int* e = indices.data() + indptr[r + 1];
          const int* q = std::lower_bound(b, e, c);
          kktloc[k] = (q != e && *q == c) ? (int) (q - indices.data()) : -1;
        }
      });
    for (auto& t: pool) t.join();
  }
}

enum { FL_OBJ = 1, FL_CON = 2, FL_JAC = 4, FL_HESS = 8, FL_GRAD = 16 };

static inline void add_to(double* p, double v, bool atomic) {
  if (!atomic) {
    *p += v;
    return;
  }
  std::atomic_ref<double> r(*p);
  r.fetch_add(v, std::memory_order_relaxed);
}

void Problem::eval(int threads, int flags, double obj_scale, const double* X, const double* LE, const double* LI,
                   const double* solver_coeffs, double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt) const {
  const long long nnz = (long long) indices.size();
  if (kkt) std::fill_n(kkt, nnz, 0.0);  // PSIOPT::evalNLP zeroes the value array, PSIOPT.cpp:107
  if (pgx) std::fill_n(pgx, nvars, 0.0);
  if (agx) std::fill_n(agx, nvars, 0.0);
  if (fxe) std::fill_n(fxe, neq, 0.0);
  if (fxi) std::fill_n(fxi, niq, 0.0);
  threads = std::max(1, threads);
  std::vector<double> vals(threads, 0.0);
  const bool atomic = threads > 1;
  auto work = [&](int thr) {
    Vec x, l, fx;
    Mat jx, hx;
    for (int fi: order) {
      const FuncEntry& f = funcs[fi];
      if (f.role == 0 && !(flags & FL_OBJ)) continue;
      if (f.role != 0 && !(flags & FL_CON)) continue;
      const int ir = f.fn->ir, orr = f.fn->orows;
      // SolverIndexingData::thread_split, IndexingData.h:117-146: contiguous chunks, remainder to the first threads
      const int per = f.nappl / threads, rem = f.nappl % threads;
      const int begin = thr * per + std::min(thr, rem), end = begin + per + (thr < rem ? 1 : 0);
      const bool dohess = (flags & FL_HESS) && !f.linear;
      const bool dojac = (flags & FL_JAC) && f.role != 0;
      const bool dograd = (flags & FL_GRAD);
      const int want = (flags & FL_HESS) ? 2 : ((dojac || dograd) ? 1 : 0);
      const double* L = (f.role == 1) ? LE : LI;
      double* FX = (f.role == 1) ? fxe : fxi;
      x.resize(ir);
      l.resize(orr);
      fx.resize(orr);
      for (int a = begin; a < end; a++) {
        const int* v = &f.V[(size_t) a * ir];
        const int* c = f.role ? &f.C[(size_t) a * orr] : nullptr;
        for (int i = 0; i < ir; i++) x[i] = X[v[i]];  // gatherInput, ComputableBase.h:351-361
        for (int i = 0; i < orr; i++) l[i] = (f.role == 0) ? obj_scale : (L ? L[c[i]] : 0.0);
        f.fn->compute(x.data(), l.data(), f.nappconst ? &f.appconst[(size_t) a * f.nappconst] : nullptr, fx.data(), jx, hx, want);
        if (f.role == 0) {
          vals[thr] += fx[0] * obj_scale;  // DenseScalarFunctionBase.h:48-79
          if (dograd && pgx)
            for (int i = 0; i < ir; i++) add_to(&pgx[v[i]], jx(0, i) * obj_scale, atomic);
        } else {
          if (FX)
            for (int i = 0; i < orr; i++) add_to(&FX[c[i]], fx[i], atomic);
          if (dograd && agx)
            for (int i = 0; i < ir; i++) {
              double s = 0;
              for (int j = 0; j < orr; j++) s += jx(j, i) * l[j];
              add_to(&agx[v[i]], s, atomic);
            }
        }
        if (kkt && (flags & (FL_JAC | FL_HESS))) {  // KKTFillAll, DenseFunctionBase.h:1413-1466
          long long k = trip_base[fi] + (long long) a * f.triplets();
          for (int i = 0; i < ir; i++) {
            if (!f.linear) {
              if (dohess)
                for (int j = i; j < ir; j++) add_to(&kkt[kktloc[k + (j - i)]], hx(j, i), atomic);
              k += ir - i;
            }
            if (f.role != 0) {
              if (dojac)
                for (int j = 0; j < orr; j++) add_to(&kkt[kktloc[k + j]], jx(j, i), atomic);
              k += orr;
            }
          }
        }
      }
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads - 1; t++) th.emplace_back(work, t);
    work(threads - 1);
    for (auto& t: th) t.join();
  }
  if (val) {
    *val = 0;
    for (double v: vals) *val += v;
  }
  if (kkt && solver_coeffs)  // fillSolverCoeffs, NonLinearProgram.h:264-290
    for (long long i = 0; i < n_solver; i++) kkt[kktloc[(size_t) (n_user + i)]] += solver_coeffs[i];
}

}  // namespace orc

// ------------------------------------------------------------------------------------------------
// C interface for ctypes (tests / bench baseline only)
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;
#define ORC_TRY try {
#define ORC_CATCH                   \
  }                                 \
  catch (const std::exception& e) { \
    g_err = e.what();               \
    return -1;                      \
  }

extern "C" {
const char* orc_last_error(void) { return g_err.c_str(); }
void* orc_problem_create(int nvars, int neq, int niq) {
  auto* p = new orc::Problem();
  p->nvars = nvars;
  p->neq = neq;
  p->niq = niq;
  return p;
}
void orc_problem_destroy(void* p) { delete (orc::Problem*) p; }
int orc_problem_add_function(void* pp, int role, const char* desc, int linear, int nappl, const int* vindex, const int* cindex,
                             const double* appconst, int nappconst) {
  ORC_TRY
  auto* p = (orc::Problem*) pp;
  orc::FuncEntry e;
  e.role = role;
  e.linear = linear != 0;
  e.nappl = nappl;
  e.nappconst = nappconst;
  e.fn = std::shared_ptr<orc::Function>(orc::parse_function(desc).release());
  e.V.assign(vindex, vindex + (size_t) nappl * e.fn->ir);
  if (role != 0) e.C.assign(cindex, cindex + (size_t) nappl * e.fn->orows);
  if (nappconst > 0) e.appconst.assign(appconst, appconst + (size_t) nappl * nappconst);
  p->funcs.push_back(std::move(e));
  return (int) p->funcs.size() - 1;
  ORC_CATCH
}
int orc_problem_analyze(void* pp) {
  ORC_TRY((orc::Problem*) pp)->analyze();
  return 0;
  ORC_CATCH
}
// out = {kktdim, nnz, n_user, n_solver}
int orc_problem_dims(void* pp, long long out[4]) {
  auto* p = (orc::Problem*) pp;
  out[0] = p->kktdim;
  out[1] = (long long) p->indices.size();
  out[2] = p->n_user;
  out[3] = p->n_solver;
  return 0;
}
int orc_problem_pattern(void* pp, int* indptr, int* indices, int* kktloc) {
  auto* p = (orc::Problem*) pp;
  if (indptr) std::memcpy(indptr, p->indptr.data(), p->indptr.size() * sizeof(int));
  if (indices) std::memcpy(indices, p->indices.data(), p->indices.size() * sizeof(int));
  if (kktloc) std::memcpy(kktloc, p->kktloc.data(), p->kktloc.size() * sizeof(int));
  return 0;
}
int orc_problem_eval(void* pp, int threads, int flags, double obj_scale, const double* X, const double* LE, const double* LI,
                     const double* solver_coeffs, double* val, double* pgx, double* agx, double* fxe, double* fxi, double* kkt) {
  ORC_TRY((orc::Problem*) pp)->eval(threads, flags, obj_scale, X, LE, LI, solver_coeffs, val, pgx, agx, fxe, fxi, kkt);
  return 0;
  ORC_CATCH
}
// One application of one function: fx [or], jx [or][ir], gx [ir] = jx^T l, hx [ir][ir].  Mirrors computeall() in the
// reference's derivative tests (asset_asrl/test/test_VectorFunctions/__init__.py:11-67).
int orc_function_computeall(const char* desc, const double* x, const double* l, const double* appconst, double* fx, double* jx,
                            double* gx, double* hx, int* ir_out, int* or_out) {
  ORC_TRY
  auto f = orc::parse_function(desc);
  if (ir_out) *ir_out = f->ir;
  if (or_out) *or_out = f->orows;
  if (!x) return 0;
  orc::Mat J, H;
  std::vector<double> F(f->orows);
  f->compute(x, l, appconst, F.data(), J, H, 2);
  if (fx) std::memcpy(fx, F.data(), F.size() * sizeof(double));
  if (jx) std::memcpy(jx, J.a.data(), J.a.size() * sizeof(double));
  if (hx) std::memcpy(hx, H.a.data(), H.a.size() * sizeof(double));
  if (gx) {
    auto g = orc::atv(J, l);
    std::memcpy(gx, g.data(), g.size() * sizeof(double));
  }
  return 0;
  ORC_CATCH
}
int orc_hardware_threads(void) { return (int) std::thread::hardware_concurrency(); }
}
