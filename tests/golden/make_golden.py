"""Generates tests/golden/*.json: 50-digit mpmath evaluations of the collocation functions on the hot path.

The reference binary cannot be built in this image and its test-suite holds no per-iterate
vectors, so these fixtures are an INDEPENDENT high-precision statement of the defining formulas:
  * values are the collocation formulas of LGLDefects.h:57-122 / TrapezoidalDefects.h:146-183 /
    LGLIntegrals.h:9-73 / LGLControlSplines.h:88-113 / MeshSpacingConstraints.h:120-128 written in mpmath with the
    reference's decimal coefficient literals (LGLCoeffs.h);
  * Jacobians and lambda-weighted Hessians are mpmath.diff (high-order finite differences at 50 digits) of those
    values -- no hand-derived derivative formula enters the fixture.
Run:  python tests/golden/make_golden.py        (takes a few minutes; the JSON files are committed)
"""
import json
import os
import sys

import mpmath as mp
import numpy as np

mp.mp.dps = 50
HERE = os.path.dirname(os.path.abspath(__file__))
M = mp.mpf


def coeffs(CS):
    c = {}
    if CS == 2:
        c["inter"] = [M("0.5")]
        c["card"] = [M(0), M(1)]
        c["XI"] = [[M("0.5"), M("0.5")]]
        c["UP"] = [[M("0.5"), M("0.5")]]
        c["DXI"] = [[M(1) / 8, -M(1) / 8]]
        c["XD"] = [[M(1), M(-1)]]
        c["DXD"] = [[M(1) / 6, M(1) / 6]]
        c["IDXD"] = [M(4) / 6]
        c["RW"] = [M("0.5"), M("0.5")]
    elif CS == 3:
        s = M("4.58257569495584")
        q, w = M(1) / 686, M(1) / 360
        c["inter"] = [M("0.172673164646011"), M("0.827326835353989")]
        c["card"] = [M(0), M("0.5"), M(1)]
        c["XI"] = [[(39 * s + 231) * q, 224 * q, (-39 * s + 231) * q], [(-39 * s + 231) * q, 224 * q, (39 * s + 231) * q]]
        c["DXI"] = [[(3 * s + 21) * q, -16 * s * q, (3 * s - 21) * q], [(-3 * s + 21) * q, 16 * s * q, (-3 * s - 21) * q]]
        c["XD"] = [[(32 * s + 180) * w, -64 * s * w, (32 * s - 180) * w], [(-32 * s + 180) * w, 64 * s * w, (-32 * s - 180) * w]]
        c["DXD"] = [[(9 + s) * w, 64 * w, (9 - s) * w], [(9 - s) * w, 64 * w, (9 + s) * w]]
        c["UP"] = [[M("0.541612549639704"), M("0.571428571428571"), M("-0.113041121068274")],
                   [M("-0.113041121068274"), M("0.571428571428571"), M("0.541612549639704")]]
        c["IDXD"] = [98 * w, 98 * w]
        c["RW"] = [M(1) / 6, M(2) / 3, M(1) / 6]
        c["UZ"] = [[M(-3), M(4), M(-1)]]
        c["UO"] = [[M(1), M(-4), M(3)]]
    else:
        c["card"] = [M(0), M("2.65575603264643e-1"), M("7.34424396735357e-1"), M(1)]
        c["inter"] = [M("8.48880518607166e-2"), M("0.5"), M("9.15111948139283e-1")]
        S = lambda rows: [[M(v) for v in r] for r in rows]  # noqa: E731
        c["XI"] = S([["6.18612232711785e-1", "3.34253095933642e-1", "1.52679626438851e-2", "3.18667087106879e-2"],
                     ["1.41445282326366e-1", "3.58554717673634e-1", "3.58554717673634e-1", "1.41445282326366e-1"],
                     ["3.18667087106879e-2", "1.52679626438851e-2", "3.34253095933642e-1", "6.18612232711785e-1"]])
        c["DXI"] = S([["2.57387738427162e-2", "-5.50098654524528e-2", "-1.53026046503702e-2", "-2.38759243962924e-3"],
                      ["9.92317607754556e-3", "9.62835932121973e-2", "-9.62835932121973e-2", "-9.92317607754556e-3"],
                      ["2.38759243962924e-3", "1.53026046503702e-2", "5.50098654524528e-2", "-2.57387738427162e-2"]])
        c["UP"] = S([["0.550643660407289", "0.551767574740443", "-0.153490305524281", "0.0510790703765507"],
                     ["-0.140877081724073", "0.640877081724073", "0.640877081724073", "-0.140877081724073"],
                     ["0.0510790703765507", "-0.153490305524281", "0.551767574740443", "0.550643660407289"]])
        c["XD"] = S([["8.84260109348311e-1", "-8.23622559094327e-1", "-2.35465327970606e-2", "-3.70910174569208e-2"],
                     ["7.86488731947674e-2", "8.00076026297266e-1", "-8.00076026297266e-1", "-7.86488731947674e-2"],
                     ["3.70910174569208e-2", "2.35465327970606e-2", "8.23622559094327e-1", "-8.84260109348311e-1"]])
        c["DXD"] = S([["1.62213410652341e-2", "9.71662045547156e-2", "1.85682012187242e-2", "2.74945307600086e-3"],
                      ["4.83872966828888e-3", "1.00138284831491e-1", "1.00138284831491e-1", "4.83872966828888e-3"],
                      ["2.74945307600086e-3", "1.85682012187242e-2", "9.71662045547156e-2", "1.62213410652341e-2"]])
        c["IDXD"] = [M("1.38413023680783e-1"), M("2.43809523809524e-1"), M("1.38413023680783e-1")]
        a3 = [M("-5.12701665379258"), M("10.9353308042859"), M("-10.9353308042859"), M("5.12701665379258")]
        a2 = [M("10.2540333075852"), M("-18.9665045333251"), M("13.8394878795326"), M("-5.12701665379258")]
        a1 = [M("-6.12701665379258"), M("8.03117372903925"), M("-2.90415707524666"), M(1)]
        a0 = [M(1), M(0), M(0), M(0)]
        # The reference evaluates these literal expressions in double precision (static constexpr double); the fixture
        # rounds each product to double first so that only the *formula*, not the coefficient rounding, is compared.
        D = lambda v: M(float(v))  # noqa: E731
        c["RW"] = [D(D(a3[j]) / 4 + D(a2[j]) / 3 + D(a1[j]) / 2 + a0[j]) for j in range(4)]
        c["UZ"] = [[D(a1[j]) for j in range(4)], [D(D(a2[j]) * 2) for j in range(4)]]
        c["UO"] = [[D(D(a3[j]) * 3 + D(a2[j]) * 2 + D(a1[j])) for j in range(4)], [D(D(a3[j]) * 6 + D(a2[j]) * 2) for j in range(4)]]
    # coefficients are doubles in the reference: round every one to double precision
    for k, v in c.items():
        c[k] = [[M(float(e)) for e in r] for r in v] if isinstance(v[0], list) else [M(float(e)) for e in v]
    return c


# ---------------------------------------------------------------- ODEs (the reference's example dynamics)
def ode_lowthrust(X):  # examples/SimpleLowThrust.py:32-39, mu = 1, acc = 0.02
    r, v, u = X[0:3], X[3:6], X[7:10]
    r3 = mp.sqrt(r[0] ** 2 + r[1] ** 2 + r[2] ** 2) ** 3
    return [v[0], v[1], v[2]] + [-r[k] / r3 + M(float(0.02)) * u[k] for k in range(3)]


def ode_brach(X):  # examples/Brachistochrone.py:23-31
    v, th = X[2], X[4]
    g = M(float(9.81))
    return [mp.sin(th) * v, -mp.cos(th) * v, g * mp.cos(th)]


def ode_vdp_param(X):  # VanDerPol with mu as an ODE parameter (exercises PV > 0): x, xd | t | u | mu
    x0, x1, u, mu = X[0], X[1], X[3], X[4]
    return [x1, (1 - x0 * x0) * x1 * mu - x0 + u]


ODES = {"lowthrust": (ode_lowthrust, 6, 3, 0), "brachistochrone": (ode_brach, 3, 1, 0), "vanderpol_mu": (ode_vdp_param, 2, 1, 1)}


def lgl_defect(ode, XV, UV, PV, CS):
    c = coeffs(CS)
    n = XV + 1 + UV

    def f(*x):
        p = list(x[CS * n:])
        CX = [list(x[j * n:(j + 1) * n]) + p for j in range(CS)]
        CF = [ode(X) for X in CX]
        h = CX[CS - 1][XV] - CX[0][XV]
        out = []
        for i in range(CS - 1):
            IX = [M(0)] * (n + PV)
            IX[XV] = CX[0][XV] + h * c["inter"][i]
            for k in range(PV):
                IX[n + k] = p[k]
            d = [M(0)] * XV
            for j in range(CS):
                for k in range(XV):
                    IX[k] += c["XI"][i][j] * CX[j][k] + c["DXI"][i][j] * h * CF[j][k]
                    d[k] += c["XD"][i][j] * CX[j][k] + c["DXD"][i][j] * h * CF[j][k]
                for k in range(UV):
                    IX[XV + 1 + k] += c["UP"][i][j] * CX[j][XV + 1 + k]
            IF = ode(IX)
            out += [d[k] + h * c["IDXD"][i] * IF[k] for k in range(XV)]
        return out
    return f, CS * n + PV, (CS - 1) * XV


def trap_defect(ode, XV, UV, PV):
    n = XV + 1 + UV

    def f(*x):
        p = list(x[2 * n:])
        X0, X1 = list(x[0:n]) + p, list(x[n:2 * n]) + p
        F0, F1 = ode(X0), ode(X1)
        h = X1[XV] - X0[XV]
        return [-((X1[k] - X0[k]) - h / 2 * (F0[k] + F1[k])) for k in range(XV)]
    return f, 2 * n + PV, XV


def control_spline(CS, UV, order):
    c = coeffs(CS)
    tu = UV + 1

    def f(*x):
        t = lambda i: x[i * tu]  # noqa: E731
        u = lambda i, k: x[i * tu + 1 + k]  # noqa: E731
        h0, h1 = t(CS - 1) - t(0), t(2 * CS - 2) - t(CS - 1)
        out = []
        for j in range(order):
            for k in range(UV):
                out.append(sum(c["UO"][j][i] * u(i, k) / h0 ** (j + 1) - c["UZ"][j][i] * u(i + CS - 1, k) / h1 ** (j + 1)
                               for i in range(CS)))
        return out
    return f, (2 * CS - 1) * tu, UV * order


def mesh_spacing(CS):
    c = coeffs(CS)

    def f(*x):
        h = x[CS - 1] - x[0]
        return [c["card"][i + 1] - (x[1 + i] - x[0]) / h for i in range(CS - 2)]
    return f, CS, CS - 2


def integral_unorm(CS, xv):  # h * sum_i W_i * (0.5 * |x_i|^2)   (examples' control-effort objective)
    c = coeffs(CS)
    xt = xv + 1

    def f(*x):
        h = x[(CS - 1) * xt + xv] - x[xv]
        return [h * sum(c["RW"][i] * M("0.5") * sum(x[i * xt + k] ** 2 for k in range(xv)) for i in range(CS))]
    return f, CS * xt, 1


def derivatives(f, x, lam):
    nin = len(x)
    fx = f(*x)
    J = [[None] * nin for _ in fx]
    for i in range(nin):
        order = tuple(1 if k == i else 0 for k in range(nin))
        for r in range(len(fx)):
            J[r][i] = mp.diff(lambda *a, r=r: f(*a)[r], tuple(x), order)
    g = lambda *a: sum(l * v for l, v in zip(lam, f(*a)))  # noqa: E731
    H = [[None] * nin for _ in range(nin)]
    for i in range(nin):
        for j in range(i, nin):
            order = [0] * nin
            order[i] += 1
            order[j] += 1
            H[i][j] = H[j][i] = mp.diff(g, tuple(x), tuple(order))
    return fx, J, H


def fl(v):
    if isinstance(v, list):
        return [fl(e) for e in v]
    return float(v)


def node_states(name, S, rng):
    if name == "lowthrust":
        th = np.linspace(0.3, 0.9, S)
        T = np.zeros((S, 10))
        T[:, 0], T[:, 1], T[:, 2] = np.cos(th), np.sin(th), 0.1
        T[:, 3], T[:, 4], T[:, 5] = -np.sin(th), np.cos(th), 0.05
        T[:, 6] = th
        T[:, 7:] = 0.5
        T[:, :6] += 0.01 * rng.uniform(-1, 1, (S, 6))
        T[:, 7:] += 0.1 * rng.uniform(-1, 1, (S, 3))
    elif name == "brachistochrone":
        t = np.linspace(0.1, 0.5, S)
        T = np.stack([t, 1 - t, 1 + 3 * t, t, 0.3 + t], axis=1)
        T[:, [0, 1, 2, 4]] += 0.01 * rng.uniform(-1, 1, (S, 4))
    else:
        t = np.linspace(0.2, 1.1, S)
        T = np.stack([np.cos(t), -np.sin(t), t, 0.1 * np.sin(2 * t)], axis=1)
        T[:, [0, 1, 3]] += 0.01 * rng.uniform(-1, 1, (S, 3))
    return T


def main():
    rng = np.random.default_rng(20260104)
    cases = []
    for ode_name in ("lowthrust", "brachistochrone", "vanderpol_mu"):
        ode, XV, UV, PV = ODES[ode_name]
        for mode, CS in (("LGL3", 2), ("LGL5", 3), ("LGL7", 4), ("Trapezoidal", 2)):
            if ode_name == "lowthrust" and mode == "LGL7" and "--full" not in sys.argv:
                pass  # kept: the biggest case (IR = 40) takes about a minute
            f, nin, nout = (trap_defect(ode, XV, UV, PV) if mode == "Trapezoidal" else lgl_defect(ode, XV, UV, PV, CS))
            T = node_states(ode_name, CS, rng)
            x = list(T.ravel()) + ([0.7] if PV else [])
            lam = list(rng.uniform(-1, 1, nout))
            cases.append(dict(name=f"{ode_name}_{mode}_defect", kind="defect", ode=ode_name, mode=mode, XV=XV, UV=UV, PV=PV,
                              CS=CS, f=f, x=x, lam=lam))
    for CS, order in ((3, 1), (4, 1), (4, 2)):
        f, nin, nout = control_spline(CS, 3, order)
        ts = np.cumsum(rng.uniform(0.1, 0.3, 2 * CS - 1))
        x = []
        for i in range(2 * CS - 1):
            x += [ts[i]] + list(rng.uniform(-1, 1, 3))
        cases.append(dict(name=f"control_spline_CS{CS}_order{order}", kind="control_spline", CS=CS, UV=3, order=order, f=f, x=x,
                          lam=list(rng.uniform(-1, 1, nout))))
    for CS in (3, 4):
        f, nin, nout = mesh_spacing(CS)
        cases.append(dict(name=f"lgl_mesh_spacing_CS{CS}", kind="lgl_mesh_spacing", CS=CS, f=f, x=list(np.cumsum(rng.uniform(0.1, 0.3, CS))),
                          lam=list(rng.uniform(-1, 1, nout))))
    for CS in (2, 3, 4):
        f, nin, nout = integral_unorm(CS, 3)
        ts = np.cumsum(rng.uniform(0.1, 0.3, CS))
        x = []
        for i in range(CS):
            x += list(rng.uniform(-1, 1, 3)) + [ts[i]]
        cases.append(dict(name=f"integral_unorm_CS{CS}", kind="integral_unorm", CS=CS, xv=3, f=f, x=x, lam=[0.7]))
    out = []
    for c in cases:
        print("golden:", c["name"], "inputs", len(c["x"]), flush=True)
        x = [M(float(v)) for v in c["x"]]
        lam = [M(float(v)) for v in c["lam"]]
        fx, J, H = derivatives(c.pop("f"), x, lam)
        c.update(x=fl(x), lam=fl(lam), fx=fl(fx), jx=fl(J), hx=fl(H))
        out.append(c)
    with open(os.path.join(HERE, "collocation_functions.json"), "w") as fh:
        json.dump(out, fh)
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
