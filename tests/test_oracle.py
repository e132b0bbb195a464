"""Pins the CPU oracle (oracle/oracle.cpp) before anything is compared against it.

The reference binary is unbuildable here and its tests hold no per-iterate vectors ("parity unpinned").
Pins used: (1) the reference's own derivative self-checks -- jx^T L == gx to 1e-12, finite differences of the value to
1e-4 (asset_asrl/test/test_VectorFunctions/__init__.py:11-67, tightened here to central differences), and
(2) independent 50-digit mpmath evaluations committed in tests/golden/collocation_functions.json.
"""
import json
import os

import numpy as np
import pytest

import oracle
from helpers import expression_zoo, interp_zoo, matrix_zoo, rel_err, vf, oc, TOL
from asset_asrl_b200 import workloads as W

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "collocation_functions.json")))


def fd_check(desc, x, lam, appc=None, jtol=1e-6, htol=1e-5):
    fx, jx, gx, hx = oracle.computeall(desc, x, lam, appc)
    assert np.abs(jx.T @ lam - gx).max() < 1e-12  # reference check (i)
    n = x.size
    eps = 1e-6
    jfd = np.zeros_like(jx)
    hfd = np.zeros_like(hx)
    for i in range(n):
        xp, xm = x.copy(), x.copy()
        xp[i] += eps
        xm[i] -= eps
        fp, _, gp, _ = oracle.computeall(desc, xp, lam, appc)
        fm, _, gm, _ = oracle.computeall(desc, xm, lam, appc)
        jfd[:, i] = (fp - fm) / (2 * eps)
        hfd[:, i] = (gp - gm) / (2 * eps)
    hfd = 0.5 * (hfd + hfd.T)
    assert rel_err(jx, jfd) < jtol if np.abs(jfd).max() > 1e-8 else np.abs(jx).max() < 1e-8  # reference check (ii)
    if np.abs(hfd).max() > 1e-6:
        assert rel_err(hx, hfd) < htol  # reference check (iii)
    else:
        assert np.abs(hx).max() < 1e-6
    assert np.abs(hx - hx.T).max() <= 1e-12 * max(np.abs(hx).max(), 1.0)


@pytest.mark.parametrize("name,fun,gen", expression_zoo() + matrix_zoo() + interp_zoo(), ids=[z[0] for z in expression_zoo() + matrix_zoo() + interp_zoo()])
def test_expression_derivatives(name, fun, gen):
    rng = np.random.default_rng(11)
    for _ in range(3):
        fd_check("expr " + fun.describe(), gen(rng), rng.uniform(-1, 1, fun.ORows()))


def _golden_desc(c):
    if c["kind"] == "defect":
        ode = {"lowthrust": W.lowthrust_ode, "brachistochrone": W.brachistochrone_ode}.get(c["ode"])
        if ode is None:
            from helpers import vdp_param_ode
            ode = vdp_param_ode
        head = "trap_defect" if c["mode"] == "Trapezoidal" else "lgl_defect"
        return f"{head} {c['CS']} {c['XV']} {c['UV']} {c['PV']} " + ode().vf().describe()
    if c["kind"] == "control_spline":
        return f"control_spline {c['CS']} {c['UV']} {c['order']}"
    if c["kind"] == "lgl_mesh_spacing":
        return f"lgl_mesh_spacing {c['CS']}"
    if c["kind"] == "integral_unorm":
        return f"lgl_integral {c['CS']} {c['xv']} 0 " + (vf.Arguments(3).squared_norm() * 0.5).describe()
    raise KeyError(c["kind"])


@pytest.mark.parametrize("case", GOLDEN, ids=[c["name"] for c in GOLDEN])
def test_oracle_matches_mpmath_golden(case):
    x, lam = np.array(case["x"]), np.array(case["lam"])
    fx, jx, gx, hx = oracle.computeall(_golden_desc(case), x, lam)
    assert rel_err(fx, case["fx"]) < TOL
    assert rel_err(jx, case["jx"]) < TOL
    assert rel_err(hx, case["hx"]) < TOL


@pytest.mark.parametrize("case", [c for c in GOLDEN if c["kind"] == "defect"], ids=lambda c: c["name"])
def test_defect_finite_differences(case):
    fd_check(_golden_desc(case), np.array(case["x"]), np.array(case["lam"]), jtol=1e-6, htol=1e-4)


def test_threads_agree():
    """NLPTest protocol (NonLinearProgram.cpp:754-784): results equal across thread counts."""
    ph = W.make_phase("lowthrust", "LGL5", 23)
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI = W.multipliers(nlp)
    orc = oracle.Oracle.from_problem(nlp)
    a = orc.eval("KKT", 1.0, X, LE, LI, threads=1)
    for t in (2, 5):
        b = orc.eval("KKT", 1.0, X, LE, LI, threads=t)
        for k in ("KKT", "FXE", "FXI", "PGX", "AGX"):
            assert rel_err(b[k], a[k]) < 1e-13


def test_quaternion_product_known_answers():
    """FunctionQuatProduct / QuatRotation (VectorProducts.h:229-236, FreeFunctionsBuild.cpp:310-321) against the unit
    quaternion identities (i j = k, j k = i, k i = j, i i = -1; vector-first storage) and scipy's rotations."""
    from scipy.spatial.transform import Rotation
    qq = vf.quatProduct(vf.Arguments(8).head(4), vf.Arguments(8).tail(4))
    i, j, k, one = np.eye(4)
    for a, b, want in ((i, j, k), (j, k, i), (k, i, j), (i, i, -one), (j, i, -k), (one, k, k)):
        fx, _, _, _ = oracle.computeall("expr " + qq.describe(), np.concatenate([a, b]), np.zeros(4))
        assert np.array_equal(fx, want)
    rng = np.random.default_rng(5)
    rot = vf.quatRotate(vf.Arguments(7).head(4), vf.Arguments(7).tail3())
    for _ in range(4):
        q1, q2 = (q / np.linalg.norm(q) for q in rng.normal(size=(2, 4)))
        v = rng.normal(size=3)
        fx, _, _, _ = oracle.computeall("expr " + qq.describe(), np.concatenate([q1, q2]), np.zeros(4))
        want = (Rotation.from_quat(q1) * Rotation.from_quat(q2)).as_quat()
        assert min(np.abs(fx - want).max(), np.abs(fx + want).max()) < 1e-14
        fx, _, _, _ = oracle.computeall("expr " + rot.describe(), np.concatenate([q1, v]), np.zeros(3))
        assert np.abs(fx - Rotation.from_quat(q1).apply(v)).max() < 1e-14


def _eval(fun, x, lam):
    return oracle.computeall("expr " + fun.describe(), np.asarray(x, dtype=np.float64), np.asarray(lam, dtype=np.float64))


@pytest.mark.parametrize("spacing", ["even", "uneven"])
def test_interp_table_1d_known_answers(spacing):
    """InterpTable1D (InterpTable1D.h:133-262): the five-point slopes are exact for polynomials up to degree four and a
    cubic Hermite segment reproduces a cubic, so data sampled from a cubic come back exactly, with both derivatives;
    general data agree with scipy's CubicHermiteSpline given the same slopes, linear tables with np.interp; the central
    stencil on an even axis is the textbook (1, -8, 0, 8, -1) / 12h."""
    from scipy.interpolate import CubicHermiteSpline
    ts = np.linspace(-1.0, 2.0, 13) if spacing == "even" else np.array([-1.0, -0.7, -0.55, -0.1, 0.0, 0.3, 0.35, 0.9, 1.4, 1.5, 2.0])
    p = np.polynomial.Polynomial([0.3, -1.1, 0.7, 0.45])
    tab = vf.InterpTable1D(ts, p(ts), kind="cubic")
    assert tab.teven == (spacing == "even")
    f = tab.vf()
    for t in np.random.default_rng(2).uniform(ts[0], ts[-1], 12):
        fx, jx, _, hx = _eval(f, [t], [1.0])
        assert abs(fx[0] - p(t)) < 2e-14 and abs(jx[0, 0] - p.deriv()(t)) < 2e-13 and abs(hx[0, 0] - p.deriv(2)(t)) < 3e-12
        v, dv, d2v = tab.interp_deriv2(t)  # the host evaluation of the front end against the oracle
        assert abs(v[0] - fx[0]) < 1e-14 and abs(dv[0] - jx[0, 0]) < 1e-13 and abs(d2v[0] - hx[0, 0]) < 1e-12
    g = np.sin(1.3 * ts) + 0.2 * ts * ts
    tab = vf.InterpTable1D(ts, g, kind="cubic")
    if spacing == "even":
        h = ts[1] - ts[0]
        want = (g[:-4] - 8 * g[1:-3] + 8 * g[3:-1] - g[4:]) / (12 * h)
        assert np.abs(tab.dvs_dts[0, 2:-2] - want).max() < 1e-13
    ref = CubicHermiteSpline(ts, g, tab.dvs_dts[0])
    lin = vf.InterpTable1D(ts, g, kind="linear")
    for t in np.random.default_rng(3).uniform(ts[0], ts[-1], 12):
        fx, jx, _, hx = _eval(tab.vf(), [t], [1.0])
        assert abs(fx[0] - ref(t)) < 1e-14 and abs(jx[0, 0] - ref.derivative()(t)) < 1e-13 and abs(hx[0, 0] - ref.derivative(2)(t)) < 1e-11
        fx, jx, _, hx = _eval(lin.vf(), [t], [1.0])
        assert abs(fx[0] - np.interp(t, ts, g)) < 1e-14 and hx[0, 0] == 0.0
    # outside the table the end segments extrapolate (get_telem clamps, :193-196)
    fx, _, _, _ = _eval(lin.vf(), [ts[-1] + 0.25], [1.0])
    assert abs(fx[0] - (g[-1] + 0.25 * (g[-1] - g[-2]) / (ts[-1] - ts[-2]))) < 1e-14


@pytest.mark.parametrize("spacing", ["even", "uneven"])
def test_interp_table_2d_known_answers(spacing):
    """InterpTable2D (InterpTable2D.h:122-390): data sampled from a polynomial of degree three in x and in y come back
    exactly with gradient and Hessian; general data agree with two passes of scipy's CubicHermiteSpline over the same
    node derivatives (a bicubic Hermite patch is the tensor product of the one-dimensional ones); the linear kind is
    bilinear interpolation."""
    from scipy.interpolate import CubicHermiteSpline
    if spacing == "even":
        xs, ys = np.linspace(0.0, 1.8, 10), np.linspace(-1.0, 3.0, 9)
    else:
        xs = np.array([0.0, 0.2, 0.4, 0.6, 0.8, 1.0, 1.2, 1.4, 1.6, 1.8]) ** 1.5
        ys = np.array([-0.5, 0.0, 5.0, 10.0, 15.0, 20.0, 25.0, 30.0, 40.0, 50.0, 70.0]) / 20.0
    px, py = np.polynomial.Polynomial([0.2, 1.0, -0.6, 0.3]), np.polynomial.Polynomial([-0.4, 0.5, 0.25, -0.1])
    tab = vf.InterpTable2D(xs, ys, py(ys)[:, None] * px(xs)[None, :], kind="cubic")
    assert tab.xeven == (spacing == "even") and tab.yeven == (spacing == "even")
    rng = np.random.default_rng(4)
    for _ in range(10):
        x, y = rng.uniform(xs[0], xs[-1]), rng.uniform(ys[0], ys[-1])
        fx, jx, _, hx = _eval(tab.vf(), [x, y], [1.0])
        assert abs(fx[0] - px(x) * py(y)) < 1e-13
        assert np.abs(jx[0] - [px.deriv()(x) * py(y), px(x) * py.deriv()(y)]).max() < 1e-12
        want = np.array([[px.deriv(2)(x) * py(y), px.deriv()(x) * py.deriv()(y)], [px.deriv()(x) * py.deriv()(y), px(x) * py.deriv(2)(y)]])
        assert np.abs(hx - want).max() < 1e-10
    Z = np.sin(xs[None, :] + 0.3 * ys[:, None]) * (1.0 + 0.1 * ys[:, None] ** 2) + xs[None, :] ** 2
    tab = vf.InterpTable2D(xs, ys, Z, kind="cubic")
    lin = vf.InterpTable2D(xs, ys, Z, kind="linear")
    for _ in range(10):
        x, y = rng.uniform(xs[0], xs[-1]), rng.uniform(ys[0], ys[-1])
        # along x at every y node: value and y-slope; then along y
        vals = np.array([CubicHermiteSpline(xs, Z[j], tab.dzxs[j])(x) for j in range(ys.size)])
        slopes = np.array([CubicHermiteSpline(xs, tab.dzys[j], tab.dzys_dxs[j])(x) for j in range(ys.size)])
        fx, _, _, _ = _eval(tab.vf(), [x, y], [1.0])
        assert abs(fx[0] - CubicHermiteSpline(ys, vals, slopes)(y)) < 1e-13
        fx, _, _, _ = _eval(lin.vf(), [x, y], [1.0])
        row = np.array([np.interp(x, xs, Z[j]) for j in range(ys.size)])
        assert abs(fx[0] - np.interp(y, ys, row)) < 1e-14


def test_matrix_functions_known_answers():
    """Matrix views, products and inverses (MatrixFunctionBuild.cpp:17-271, MatrixProduct.h, MatrixInverse.h) against numpy,
    in both storage orders; the inverse's Jacobian against -kron(A^-T, A^-1)."""
    rng = np.random.default_rng(8)
    x = vf.Arguments(12)
    a, b = rng.normal(size=(2, 3)), rng.normal(size=(3, 2))
    xin = np.concatenate([a.T.ravel(), b.ravel()])  # A column-major, B row-major
    A_, B_ = vf.ColMatrix(x.head(6), 2, 3), vf.RowMatrix(x.tail(6), 3, 2)
    fx, _, _, _ = _eval((A_ * B_).vf(), xin, np.zeros(4))
    assert np.abs(fx.reshape(2, 2).T - a @ b).max() < 1e-14  # products come back column-major
    fx, _, _, _ = _eval((B_ * A_).vf(), xin, np.zeros(9))
    assert np.abs(fx.reshape(3, 3).T - b @ a).max() < 1e-14
    fx, _, _, _ = _eval((A_.transpose() * B_.transpose()).vf(), xin, np.zeros(9))
    assert np.abs(fx.reshape(3, 3).T - a.T @ b.T).max() < 1e-14
    v = np.array([0.3, -1.2, 2.0])
    fx, _, _, _ = _eval(vf.matmul(A_, v), xin, np.zeros(2))
    assert np.abs(fx - a @ v).max() < 1e-14
    for n in (2, 3, 4, 6):
        m = rng.normal(size=(n, n)) + np.eye(n)[rng.permutation(n)] * 3.0  # needs row exchanges
        xm = vf.Arguments(n * n)
        for cls, flat, back in ((vf.ColMatrix, m.T.ravel(), lambda f: f.reshape(n, n).T), (vf.RowMatrix, m.ravel(), lambda f: f.reshape(n, n))):
            fx, jx, _, _ = _eval(cls(xm, n, n).inverse().vf(), flat, np.zeros(n * n))
            inv = np.linalg.inv(m)
            assert np.abs(back(fx) - inv).max() < 1e-12 * np.abs(inv).max()
            if cls is vf.ColMatrix:  # d vec(A^-1) / d vec(A) = -(A^-T kron A^-1) for column-major vec
                assert np.abs(jx + np.kron(inv.T, inv)).max() < 1e-11 * np.abs(inv).max() ** 2


@pytest.mark.parametrize("D", [3, 4])
def test_interp_table_nd_known_answers(D):
    """InterpTable3D / 4D: data sampled from a product of cubics come back exactly with gradient and Hessian (the
    tri- / quadri-cubic Hermite patch reproduces them); the linear kind agrees with scipy's RegularGridInterpolator."""
    from scipy.interpolate import RegularGridInterpolator
    axes = [np.linspace(0.0, 1.5, 6), np.array([-1.0, -0.6, -0.1, 0.2, 0.9, 1.3, 2.0]), np.linspace(2.0, 3.0, 5),
            np.array([0.0, 0.3, 0.5, 1.0, 1.2, 1.9])][:D]
    polys = [np.polynomial.Polynomial(c) for c in ([0.2, 1.0, -0.6, 0.3], [-0.4, 0.5, 0.25, -0.1], [1.0, -0.3, 0.2, 0.05], [0.7, 0.1, -0.2, 0.15])][:D]
    grid = np.meshgrid(*axes, indexing="ij")
    data = np.ones_like(grid[0])
    for g, p in zip(grid, polys):
        data = data * p(g)
    Tab = vf.InterpTable3D if D == 3 else vf.InterpTable4D
    tab = Tab(*axes, data, kind="cubic")
    rng = np.random.default_rng(6)
    lo, hi = np.array([a[0] for a in axes]), np.array([a[-1] for a in axes])
    for _ in range(8):
        x = rng.uniform(lo, hi)
        fx, jx, _, hx = _eval(tab.vf(), x, [1.0])
        vals = np.array([p(v) for p, v in zip(polys, x)])
        d1 = np.array([p.deriv()(v) for p, v in zip(polys, x)])
        d2 = np.array([p.deriv(2)(v) for p, v in zip(polys, x)])
        assert abs(fx[0] - vals.prod()) < 1e-12
        for a in range(D):
            ga = np.prod([d1[d] if d == a else vals[d] for d in range(D)])
            assert abs(jx[0, a] - ga) < 1e-11
            for b in range(D):
                hab = np.prod([(d2[d] if a == b else d1[d]) if d in (a, b) else vals[d] for d in range(D)])
                assert abs(hx[a, b] - hab) < 1e-9
    wavy = np.sin(sum((k + 1) * 0.4 * g for k, g in enumerate(grid)))
    lin = Tab(*axes, wavy, kind="linear")
    ref = RegularGridInterpolator(axes, wavy, method="linear")
    for _ in range(8):
        x = rng.uniform(lo, hi)
        fx, _, _, _ = _eval(lin.vf(), x, [1.0])
        assert abs(fx[0] - ref(x[None, :])[0]) < 1e-13


def test_interp_tables_host_evaluation_matches_the_oracle():
    """tab(x, y[, z[, w]]) on numbers (the reference's InterpTable*::interp) against the oracle's evaluation of tab.vf()."""
    from helpers import interp_tables
    rng = np.random.default_rng(9)
    for nm, tab in interp_tables().items():
        if nm.startswith("1d"):
            continue
        if nm.startswith("2d"):
            lo, hi = np.array([tab.xs[0], tab.ys[0]]), np.array([tab.xs[-1], tab.ys[-1]])
        else:
            lo, hi = np.array([a[0] for a in tab.axes]), np.array([a[-1] for a in tab.axes])
        for _ in range(5):
            x = rng.uniform(lo, hi)
            fx, _, _, _ = _eval(tab.vf(), x, [1.0])
            assert abs(tab(*x) - fx[0]) <= 1e-13 * max(abs(fx[0]), 1.0), nm
