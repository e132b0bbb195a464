#include "builtins.hpp"

#include <stdexcept>

namespace ab {

namespace {

LGLTable make_lgl3() {
  LGLTable t {};
  t.CS = 2;
  t.tau_card[0] = 0.0;
  t.tau_card[1] = 1.0;
  t.tau_int[0] = 0.5;
  t.xi[0][0] = 0.5;
  t.xi[0][1] = 0.5;
  t.up[0][0] = 0.5;
  t.up[0][1] = 0.5;
  t.dxi[0][0] = 0.125;
  t.dxi[0][1] = -0.125;
  t.xd[0][0] = 1.0;
  t.xd[0][1] = -1.0;
  t.dxd[0][0] = 1.0 / 6.0;
  t.dxd[0][1] = 1.0 / 6.0;
  t.idxd[0] = 4.0 / 6.0;
  t.redw[0] = 0.5;
  t.redw[1] = 0.5;
  return t;
}

LGLTable make_lgl5() {
  LGLTable t {};
  t.CS = 3;
  const double S21 = 4.58257569495584;  // truncated literal, as in the reference
  const double a = 1.0 / 686.0;
  const double a1 = (39.0 * S21 + 231.0) * a, a2 = 224.0 * a, a3 = (-39.0 * S21 + 231.0) * a;
  const double a4 = (3.0 * S21 + 21.0) * a, a5 = (-16.0 * S21) * a, a6 = (3.0 * S21 - 21.0) * a;
  const double b = 1.0 / 686.0;
  const double b1 = (-39.0 * S21 + 231.0) * b, b2 = 224.0 * b, b3 = (39.0 * S21 + 231.0) * b;
  const double b4 = (-3.0 * S21 + 21.0) * b, b5 = (16.0 * S21) * b, b6 = (-3.0 * S21 - 21.0) * b;
  const double c = 1.0 / 360;
  const double c1 = (32.0 * S21 + 180.0) * c, c2 = -64.0 * S21 * c, c3 = (32.0 * S21 - 180.0) * c;
  const double c4 = (9.0 + S21) * c, c5 = 98.0 * c, c6 = 64.0 * c, c7 = (9.0 - S21) * c;
  const double d = 1.0 / 360.0;
  const double d1 = (-32.0 * S21 + 180.0) * d, d2 = 64.0 * S21 * d, d3 = (-32.0 * S21 - 180.0) * d;
  const double d4 = (9.0 - S21) * d, d5 = 98.0 * d, d6 = 64.0 * d, d7 = (9.0 + S21) * d;
  const double tc[3] = {0.0, 0.5, 1.0};
  const double ti[2] = {0.172673164646011, 0.827326835353989};
  for (int i = 0; i < 3; i++) t.tau_card[i] = tc[i];
  for (int i = 0; i < 2; i++) t.tau_int[i] = ti[i];
  const double xi[2][3] = {{a1, a2, a3}, {b1, b2, b3}};
  const double up[2][3] = {{0.541612549639704, 0.571428571428571, -0.113041121068274},
                           {-0.113041121068274, 0.571428571428571, 0.541612549639704}};
  const double dxi[2][3] = {{a4, a5, a6}, {b4, b5, b6}};
  const double xd[2][3] = {{c1, c2, c3}, {d1, d2, d3}};
  const double dxd[2][3] = {{c4, c6, c7}, {d4, d6, d7}};
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 3; j++) {
      t.xi[i][j] = xi[i][j];
      t.up[i][j] = up[i][j];
      t.dxi[i][j] = dxi[i][j];
      t.xd[i][j] = xd[i][j];
      t.dxd[i][j] = dxd[i][j];
    }
  t.idxd[0] = c5;
  t.idxd[1] = d5;
  t.redw[0] = 1.0 / 6.0;
  t.redw[1] = 2.0 / 3.0;
  t.redw[2] = 1.0 / 6.0;
  const double uz[3] = {-3.0, 4.0, -1.0}, uo[3] = {1.0, -4.0, 3.0};
  for (int j = 0; j < 3; j++) {
    t.uzero[0][j] = uz[j];
    t.uone[0][j] = uo[j];
  }
  return t;
}

LGLTable make_lgl7() {
  LGLTable t {};
  t.CS = 4;
  const double tc[4] = {+0.00000000000000, +2.65575603264643e-1, +7.34424396735357e-1, +1.00000000000000};
  const double ti[3] = {+8.48880518607166e-2, +0.50000000000000, +9.15111948139283e-1};
  for (int i = 0; i < 4; i++) t.tau_card[i] = tc[i];
  for (int i = 0; i < 3; i++) t.tau_int[i] = ti[i];
  const double xi[3][4] = {{+6.18612232711785e-1, +3.34253095933642e-1, +1.52679626438851e-2, +3.18667087106879e-2},
                           {+1.41445282326366e-1, +3.58554717673634e-1, +3.58554717673634e-1, +1.41445282326366e-1},
                           {+3.18667087106879e-2, +1.52679626438851e-2, +3.34253095933642e-1, +6.18612232711785e-1}};
  const double dxi[3][4] = {{+2.57387738427162e-2, -5.50098654524528e-2, -1.53026046503702e-2, -2.38759243962924e-3},
                            {+9.92317607754556e-3, +9.62835932121973e-2, -9.62835932121973e-2, -9.92317607754556e-3},
                            {+2.38759243962924e-3, +1.53026046503702e-2, +5.50098654524528e-2, -2.57387738427162e-2}};
  const double up[3][4] = {{0.550643660407289, 0.551767574740443, -0.153490305524281, 0.0510790703765507},
                           {-0.140877081724073, 0.640877081724073, 0.640877081724073, -0.140877081724073},
                           {0.0510790703765507, -0.153490305524281, 0.551767574740443, 0.550643660407289}};
  const double xd[3][4] = {{+8.84260109348311e-1, -8.23622559094327e-1, -2.35465327970606e-2, -3.70910174569208e-2},
                           {+7.86488731947674e-2, +8.00076026297266e-1, -8.00076026297266e-1, -7.86488731947674e-2},
                           {+3.70910174569208e-2, +2.35465327970606e-2, +8.23622559094327e-1, -8.84260109348311e-1}};
  const double dxd[3][4] = {{+1.62213410652341e-2, +9.71662045547156e-2, +1.85682012187242e-2, +2.74945307600086e-3},
                            {+4.83872966828888e-3, +1.00138284831491e-1, +1.00138284831491e-1, +4.83872966828888e-3},
                            {+2.74945307600086e-3, +1.85682012187242e-2, +9.71662045547156e-2, +1.62213410652341e-2}};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 4; j++) {
      t.xi[i][j] = xi[i][j];
      t.dxi[i][j] = dxi[i][j];
      t.up[i][j] = up[i][j];
      t.xd[i][j] = xd[i][j];
      t.dxd[i][j] = dxd[i][j];
    }
  t.idxd[0] = +1.38413023680783e-1;
  t.idxd[1] = +2.43809523809524e-1;
  t.idxd[2] = +1.38413023680783e-1;
  t.redw[0] = -5.12701665379258 / 4.0 + 10.2540333075852 / 3.0 - 6.12701665379258 / 2.0 + 1.0;
  t.redw[1] = 10.9353308042859 / 4.0 - 18.9665045333251 / 3.0 + 8.03117372903925 / 2.0;
  t.redw[2] = -10.9353308042859 / 4.0 + 13.8394878795326 / 3.0 - 2.90415707524666 / 2.0;
  t.redw[3] = 5.12701665379258 / 4.0 - 5.12701665379258 / 3.0 + 1.0 / 2.0;
  const double uz[2][4] = {{-6.12701665379258, 8.03117372903925, -2.90415707524666, 1.0},
                           {10.2540333075852 * 2.0, -18.9665045333251 * 2.0, +13.8394878795326 * 2.0, -5.12701665379258 * 2.0}};
  const double uo[2][4] = {{-5.12701665379258 * 3.0 + 10.2540333075852 * 2.0 - 6.12701665379258,
                            10.9353308042859 * 3.0 - 18.9665045333251 * 2.0 + 8.03117372903925,
                            -10.9353308042859 * 3.0 + 13.8394878795326 * 2.0 - 2.90415707524666,
                            5.12701665379258 * 3.0 - 5.12701665379258 * 2.0 + 1.0},
                           {-5.12701665379258 * 6.0 + 10.2540333075852 * 2.0, 10.9353308042859 * 6.0 - 18.9665045333251 * 2.0,
                            -10.9353308042859 * 6.0 + 13.8394878795326 * 2.0, 5.12701665379258 * 6.0 - 5.12701665379258 * 2.0}};
  for (int k = 0; k < 2; k++)
    for (int j = 0; j < 4; j++) {
      t.uzero[k][j] = uz[k][j];
      t.uone[k][j] = uo[k][j];
    }
  return t;
}

}  // namespace

const LGLTable& lgl_table(int CS) {
  static const LGLTable t3 = make_lgl3(), t5 = make_lgl5(), t7 = make_lgl7();
  switch (CS) {
    case 2: return t3;
    case 3: return t5;
    case 4: return t7;
    default: throw std::invalid_argument("lgl_table: CS must be 2, 3 or 4");
  }
}

ScalarFn make_lgl_defect(const Graph& g, int ode_root, int CS, int XV, int UV, int PV) {
  const LGLTable& T = lgl_table(CS);
  const int n = XV + 1 + UV, no = n + PV;
  if (g.at(ode_root).ir != no || g.at(ode_root).orows != XV) throw std::invalid_argument("make_lgl_defect: ODE sizes do not match XV/UV/PV");
  ScalarFn f;
  f.nin = CS * n + PV;
  f.nout = (CS - 1) * XV;
  SDag& d = f.dag;
  std::vector<int> in(f.nin);
  for (int i = 0; i < f.nin; i++) in[i] = d.input(i);
  std::vector<std::vector<int>> CX(CS), CF(CS);
  for (int j = 0; j < CS; j++) {
    CX[j].assign(in.begin() + j * n, in.begin() + (j + 1) * n);
    CX[j].insert(CX[j].end(), in.begin() + CS * n, in.end());
    CF[j] = lower(g, ode_root, CX[j], d);
  }
  const int tv = XV;
  const int h = d.sub(CX[CS - 1][tv], CX[0][tv]);
  f.outs.assign(f.nout, -1);
  for (int i = 0; i < CS - 1; i++) {
    std::vector<int> IX(no, -1);
    IX[tv] = d.add(CX[0][tv], d.mulc(h, T.tau_int[i]));
    for (int p = 0; p < PV; p++) IX[n + p] = in[CS * n + p];
    std::vector<int> def(XV, -1);
    for (int j = 0; j < CS; j++) {
      const int vh = d.mulc(h, T.dxi[i][j]);
      const int wh = d.mulc(h, T.dxd[i][j]);
      for (int k = 0; k < XV; k++) {
        int term = d.add(d.mulc(CX[j][k], T.xi[i][j]), d.mul(vh, CF[j][k]));
        IX[k] = (IX[k] < 0) ? term : d.add(IX[k], term);
        int dterm = d.add(d.mulc(CX[j][k], T.xd[i][j]), d.mul(wh, CF[j][k]));
        def[k] = (def[k] < 0) ? dterm : d.add(def[k], dterm);
      }
      for (int k = 0; k < UV; k++) {
        int term = d.mulc(CX[j][XV + 1 + k], T.up[i][j]);
        IX[XV + 1 + k] = (IX[XV + 1 + k] < 0) ? term : d.add(IX[XV + 1 + k], term);
      }
    }
    std::vector<int> IF = lower(g, ode_root, IX, d);
    const int hw = d.mulc(h, T.idxd[i]);
    for (int k = 0; k < XV; k++) f.outs[i * XV + k] = d.add(def[k], d.mul(hw, IF[k]));
  }
  return f;
}

ScalarFn make_trap_defect(const Graph& g, int ode_root, int XV, int UV, int PV) {
  const int n = XV + 1 + UV, no = n + PV;
  if (g.at(ode_root).ir != no || g.at(ode_root).orows != XV) throw std::invalid_argument("make_trap_defect: ODE sizes do not match XV/UV/PV");
  ScalarFn f;
  f.nin = 2 * n + PV;
  f.nout = XV;
  SDag& d = f.dag;
  std::vector<int> in(f.nin);
  for (int i = 0; i < f.nin; i++) in[i] = d.input(i);
  std::vector<int> X0(in.begin(), in.begin() + n), X1(in.begin() + n, in.begin() + 2 * n);
  X0.insert(X0.end(), in.begin() + 2 * n, in.end());
  X1.insert(X1.end(), in.begin() + 2 * n, in.end());
  auto F0 = lower(g, ode_root, X0, d), F1 = lower(g, ode_root, X1, d);
  const int h = d.sub(X1[XV], X0[XV]);
  const int h2 = d.mulc(h, 0.5);
  for (int k = 0; k < XV; k++) {
    int v = d.sub(d.sub(X1[k], X0[k]), d.mul(h2, d.add(F0[k], F1[k])));
    f.outs.push_back(d.neg(v));
  }
  return f;
}

ScalarFn make_lgl_mesh_spacing(int CS) {
  const LGLTable& T = lgl_table(CS);
  ScalarFn f;
  f.nin = CS;
  f.nout = CS - 2;
  SDag& d = f.dag;
  std::vector<int> x(CS);
  for (int i = 0; i < CS; i++) x[i] = d.input(i);
  const int h = d.sub(x[CS - 1], x[0]);
  for (int i = 0; i < CS - 2; i++) {
    int r = d.div(d.sub(x[1 + i], x[0]), h);
    f.outs.push_back(d.addc(d.neg(r), T.tau_card[i + 1]));
  }
  return f;
}

ScalarFn make_single_mesh_spacing() {
  ScalarFn f;
  f.nin = 3;
  f.nout = 1;
  SDag& d = f.dag;
  int x0 = d.input(0), x1 = d.input(1), x2 = d.input(2);
  int h = d.sub(x2, x0);
  f.outs.push_back(d.sub(d.mul(d.appconst(0), h), d.sub(x1, x0)));
  return f;
}

ScalarFn make_control_spline(int CS, int UV, int order) {
  const LGLTable& T = lgl_table(CS);
  if (CS < 3 || order < 1 || order > CS - 2) throw std::invalid_argument("make_control_spline: bad CS/order");
  const int tu = UV + 1, nn = 2 * CS - 1;
  ScalarFn f;
  f.nin = nn * tu;
  f.nout = UV * order;
  SDag& d = f.dag;
  std::vector<int> in(f.nin);
  for (int i = 0; i < f.nin; i++) in[i] = d.input(i);
  auto tnode = [&](int i) { return in[i * tu]; };
  auto unode = [&](int i, int k) { return in[i * tu + 1 + k]; };
  const int h0 = d.sub(tnode(CS - 1), tnode(0));
  const int h1 = d.sub(tnode(nn - 1), tnode(CS - 1));
  for (int j = 0; j < order; j++) {
    const int h0p = d.unary(OP_INV, d.unary(OP_POWC, h0, (double) (j + 1)));
    const int h1p = d.unary(OP_INV, d.unary(OP_POWC, h1, (double) (j + 1)));
    for (int k = 0; k < UV; k++) {
      int acc = -1;
      for (int i = 0; i < CS; i++) {
        int term = d.sub(d.mul(d.mulc(h0p, T.uone[j][i]), unode(i, k)), d.mul(d.mulc(h1p, T.uzero[j][i]), unode(i + CS - 1, k)));
        acc = (acc < 0) ? term : d.add(acc, term);
      }
      f.outs.push_back(acc);
    }
  }
  return f;
}

ScalarFn make_lgl_integral(const Graph& g, int integrand_root, int CS, int xv, int pv) {
  const LGLTable& T = lgl_table(CS);
  if (g.at(integrand_root).ir != xv + pv || g.at(integrand_root).orows != 1)
    throw std::invalid_argument("make_lgl_integral: integrand must map (xv+pv) -> 1");
  const int xtv = xv + 1;
  ScalarFn f;
  f.nin = CS * xtv + pv;
  f.nout = 1;
  SDag& d = f.dag;
  std::vector<int> in(f.nin);
  for (int i = 0; i < f.nin; i++) in[i] = d.input(i);
  const int h = d.sub(in[(CS - 1) * xtv + xv], in[xv]);
  int acc = -1;
  for (int i = 0; i < CS; i++) {
    std::vector<int> xi(in.begin() + i * xtv, in.begin() + i * xtv + xv);
    xi.insert(xi.end(), in.begin() + CS * xtv, in.end());
    int gi = lower(g, integrand_root, xi, d)[0];
    int term = d.mulc(gi, T.redw[i]);
    acc = (acc < 0) ? term : d.add(acc, term);
  }
  f.outs.push_back(d.mul(acc, h));
  return f;
}

}  // namespace ab
