"""Shared builders for the test-suite: a zoo of expressions and of small phases covering every region flag, transcription
and control mode the hot path supports.  Test-only."""
import numpy as np

import asset_asrl_b200 as ast
from asset_asrl_b200 import workloads as W

vf, oc = ast.VectorFunctions, ast.OptimalControl

TOL = 1e-12  # BASELINE.json north_star: 1e-12 relative (fp64) for constraint, Jacobian and Hessian values


def rel_err(a, b):
    """|a-b|_inf relative to the block norm |b|_inf (cancellation in Hessian entries needs a
    relative-to-block-norm comparison)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-300)


def expression_zoo():
    """(name, function, sample point generator) over the op table of SURVEY.md section 8a row a16."""
    A = vf.Arguments
    z = []
    x3 = A(3)
    x4 = A(4)
    x6 = A(6)
    pos = lambda n: (lambda rng: rng.uniform(0.5, 1.5, n))  # noqa: E731
    sym = lambda n: (lambda rng: rng.uniform(-0.9, 0.9, n))  # noqa: E731
    for nm in ("sin", "cos", "tan", "arcsin", "arccos", "arctan", "sinh", "cosh", "tanh", "arcsinh", "arctanh", "exp", "squared"):
        z.append((nm, getattr(vf, nm)(x3[0] * x3[1] + x3[2] * 0.3), sym(3)))
    for nm in ("sqrt", "log", "arccosh"):
        z.append((nm, getattr(vf, nm)(x3[0] * x3[1] + x3[2] + 1.0), pos(3)))
    z.append(("abs_sign", vf.stack([vf.abs(x3[0] * x3[1]), vf.sign(x3[2]) * x3[0]]), sym(3)))
    z.append(("pow", vf.stack([x3[0] ** 3, x3[1] ** 2.5, x3[2] ** -1.5, (x3[0] * x3[1]) ** 0.5]), pos(3)))
    z.append(("inverse", (x3[0] * x3[1] + x3[2]).inverse(), pos(3)))
    z.append(("norms", vf.stack([x4.norm(), x4.squared_norm(), x4.cubed_norm(), x4.inverse_norm(), x4.inverse_squared_norm(),
                                 x4.inverse_cubed_norm(), x4.inverse_four_norm()]), pos(4)))
    z.append(("normalized", vf.stack([x3.normalized(), x3.normalized_power2(), x3.normalized_power3(), x3.normalized_power4(),
                                      x3.normalized_power5()]), pos(3)))
    z.append(("dot_cross", vf.stack([x6.head3().dot(x6.tail3()), x6.head3().cross(x6.tail3())]), sym(6)))
    z.append(("cwise", vf.stack([x6.head3().cwiseProduct(x6.tail3()), x6.head3().cwiseQuotient(x6.tail3())]), pos(6)))
    z.append(("vec_scalar", vf.stack([x4.head3() * x4[3], x4.head3() / x4[3], x4[3] / x4[0]]), pos(4)))
    z.append(("arctan2", vf.arctan2(x3[0] * x3[2], x3[1] + x3[2]), pos(3)))
    z.append(("sum_diff", vf.sum([x6.head3(), x6.tail3(), x6.segment(1, 3)]) - x6.segment(2, 3) * 2.0 + np.array([1.0, 2.0, 3.0]), sym(6)))
    z.append(("nested", vf.Arguments(2).normalized_power3().eval(vf.stack([x3[0] * x3[1], x3[2] + 2.0])), pos(3)))
    z.append(("ifelse", vf.ifelse(x3[0] < x3[1], x3[0] * x3[2], vf.sin(x3[1]) * x3[2]), sym(3)))
    z.append(("ifelse_and_or", vf.ifelse(((x3[0] < x3[1]) & (x3[1] <= x3[2])) | (x3[2] > 0.5), x3.norm(), x3.squared_norm() * 2.0), sym(3)))
    z.append(("matmul", vf.matmul(np.arange(6.0).reshape(2, 3) - 2.0, vf.sin(x3)), sym(3)))
    z.append(("ioscaled", vf.ioscaled(x3.normalized_power3(), [2.0, 0.5, 3.0], [0.1, 10.0, -2.0]), pos(3)))
    z.append(("cwise_sum_elements", (vf.sin(x6).sum()) * x6[[0, 3, 5]].norm(), pos(6)))
    z.append(("two_body", vf.stack([x6.tail3(), x6.head3().normalized_power3() * (-1.0)]), pos(6)))
    x8, x7 = A(8), A(7)
    z.append(("quat_product", vf.quatProduct(x8.head(4), vf.sin(x8.tail(4))), sym(8)))
    z.append(("quat_kinematics", vf.quatProduct(x7.head(4), x7.tail3().padded_lower(1)) / 2.0, sym(7)))  # OptimalDocking.py:96
    z.append(("quat_rotate", vf.quatRotate(x7.head(4).normalized(), x7.tail3()), pos(7)))
    return z


def matrix_zoo():
    """Matrix functions (MatrixFunctionBuild.cpp): views, products in every storage-order pairing, inverses of sizes 2..4
    (the 4 x 4 one with row exchanges), ParsedInput.  Kept apart from expression_zoo like interp_zoo (GPU tests in
    tests/test_zz2_gpu_tables.py)."""
    A = vf.Arguments
    x4, x6, x9, x16 = A(4), A(6), A(9), A(16)
    z = []
    pos = lambda n: (lambda rng: rng.uniform(0.5, 1.5, n))  # noqa: E731
    sym = lambda n: (lambda rng: rng.uniform(-0.9, 0.9, n))  # noqa: E731
    dd = lambda n: (lambda rng: (np.eye(int(round(n ** 0.5))) * 2.5 + rng.uniform(-0.6, 0.6, (int(round(n ** 0.5)),) * 2)).ravel())  # noqa: E731
    Mc, Mr = vf.ColMatrix(x6, 2, 3), vf.RowMatrix(vf.sin(x6), 3, 2)
    z.append(("mat_products", vf.stack([(Mc * Mr).vf(), (Mr * Mc).vf(), (Mc.transpose() * Mr.transpose()).vf(), Mc * x6.tail3(),
                                        vf.matmul(Mr, np.array([0.5, -2.0])), (2.0 * Mc + np.arange(6.0).reshape(2, 3)).vf()]), sym(6)))
    z.append(("mat_inverse_2", vf.stack([vf.ColMatrix(x4, 2, 2).inverse().vf(), vf.RowMatrix(x4, 2, 2).inverse() * x4.head2()]), dd(4)))
    z.append(("mat_inverse_3", vf.stack([vf.ColMatrix(x9, 3, 3).inverse().vf(), vf.RowMatrix(x9, 3, 3).inverse() * x9.head3()]), dd(9)))
    z.append(("mat_inverse_4_pivoting", (vf.RowMatrix(x16, 4, 4).inverse() * vf.RowMatrix(x16, 4, 4).transpose()).vf(),
              lambda rng: (np.eye(4)[rng.permutation(4)] * 2.0 + rng.uniform(-0.4, 0.4, (4, 4))).ravel()))
    z.append(("parsed_input", vf.ParsedInput(A(3).normalized_power3(), [5, 0, 2], 6), pos(6)))
    return z


def interp_tables():
    """Interpolation tables over smooth data: evenly and unevenly spaced axes, scalar and vector values, both kinds."""
    te = np.linspace(-1.0, 2.0, 13)
    tu = np.array([-1.0, -0.7, -0.55, -0.1, 0.0, 0.3, 0.35, 0.9, 1.4, 1.5, 2.0])
    sc = lambda t: np.sin(1.3 * t) + 0.2 * t * t  # noqa: E731
    vec = lambda t: np.stack([np.cos(t), np.exp(-0.5 * t), t ** 3 - t])  # noqa: E731
    xs_e, ys_e = np.linspace(0.0, 1.8, 10), np.linspace(-1.0, 3.0, 9)
    xs_u = np.array([0.0, 0.2, 0.4, 0.6, 0.8, 1.0, 1.2, 1.4, 1.6, 1.8]) ** 1.5
    ys_u = np.array([-0.5, 0.0, 5.0, 10.0, 15.0, 20.0, 25.0, 30.0, 40.0, 50.0, 70.0]) / 20.0
    surf = lambda x, y: np.sin(x[None, :] + 0.3 * y[:, None]) * (1.0 + 0.1 * y[:, None] ** 2) + x[None, :] ** 2  # noqa: E731
    T = {}
    for knd in ("cubic", "linear"):
        T[f"1d_even_scalar_{knd}"] = vf.InterpTable1D(te, sc(te), kind=knd)
        T[f"1d_uneven_scalar_{knd}"] = vf.InterpTable1D(tu, sc(tu), kind=knd)
        T[f"1d_even_vector_{knd}"] = vf.InterpTable1D(te, vec(te), axis=1, kind=knd)
        T[f"1d_uneven_vector_rows_{knd}"] = vf.InterpTable1D(np.concatenate([vec(tu).T, tu[:, None]], axis=1), kind=knd)
        T[f"2d_even_{knd}"] = vf.InterpTable2D(xs_e, ys_e, surf(xs_e, ys_e), kind=knd)
        T[f"2d_uneven_{knd}"] = vf.InterpTable2D(xs_u, ys_u, surf(xs_u, ys_u), kind=knd)
        T[f"2d_mixed_{knd}"] = vf.InterpTable2D(xs_e, ys_u, surf(xs_e, ys_u), kind=knd)
        g3 = np.meshgrid(xs_e[:6], ys_u[:7], te[:5], indexing="ij")
        T[f"3d_mixed_{knd}"] = vf.InterpTable3D(xs_e[:6], ys_u[:7], te[:5], np.sin(g3[0] + 0.4 * g3[1]) * np.cos(0.7 * g3[2]) + 0.1 * g3[1] * g3[2],
                                                kind=knd)
        g4 = np.meshgrid(xs_u[:5], ys_e[:5], tu[:6], te[:5], indexing="ij")
        T[f"4d_mixed_{knd}"] = vf.InterpTable4D(xs_u[:5], ys_e[:5], tu[:6], te[:5],
                                                np.sin(g4[0] + 0.4 * g4[1]) * np.cos(0.7 * g4[2] - 0.2 * g4[3]) + 0.1 * g4[1] * g4[3], kind=knd)
    return T


def interp_zoo():
    """Expression zoo entries over the tables (kept apart from expression_zoo so that the GPU tests of the table ops
    live in a file of their own, tests/test_zz2_gpu_tables.py)."""
    A = vf.Arguments
    x3 = A(3)
    z = []
    for nm, tab in interp_tables().items():
        if nm.startswith("1d"):
            lo, hi = tab.ts[0], tab.ts[-1]
            f = tab(x3[0] * x3[1] + x3[2])  # coordinate = a nonlinear function of the inputs: chain rule through the table
            if tab.vlen == 1:
                f = vf.stack([f * x3[1], vf.sin(f)])

            def gen(rng, lo=lo, hi=hi):
                while True:
                    x = rng.uniform(-1.2, 1.2, 3)
                    if lo + 0.01 < x[0] * x[1] + x[2] < hi - 0.01:
                        return x
        elif nm[0] in "34":
            D = tab.D
            xa = A(D)
            lo = np.array([c[0] for c in tab.axes]) + 0.01
            hi = np.array([c[-1] for c in tab.axes]) - 0.01
            f = tab(*[xa[d] for d in range(D)]) * xa[0] + vf.sin(tab(xa))

            def gen(rng, lo=lo, hi=hi):
                return rng.uniform(lo, hi)
        else:
            xl, xh, yl, yh = tab.xs[0], tab.xs[-1], tab.ys[0], tab.ys[-1]
            f = tab(x3[0] * x3[2], x3[1] + x3[2]) * x3[0]

            def gen(rng, xl=xl, xh=xh, yl=yl, yh=yh):
                while True:
                    x = np.array([rng.uniform(0.3, 1.5), rng.uniform(yl, yh), rng.uniform(0.0, 1.2)])
                    if xl + 0.01 < x[0] * x[2] < xh - 0.01 and yl + 0.01 < x[1] + x[2] < yh - 0.01:
                        return x
        z.append(("interp_" + nm, f, gen))
    return z


def table_phase_zoo():
    """Phases whose ODE reads interpolation tables (kept apart from phase_zoo for the same reason as interp_zoo)."""
    return {
        "minclimb_LGL5_highest_order_spline": lambda: W.make_phase("minclimb", "LGL5", 9, control_mode="HighestOrderSpline"),
        "minclimb_LGL3": lambda: W.make_phase("minclimb", "LGL3", 8),
        "minclimb_LGL7_block_constant": lambda: W.make_phase("minclimb", "LGL7", 5, control_mode="BlockConstant"),
        "minclimb_Trapezoidal": lambda: W.make_phase("minclimb", "Trapezoidal", 10),
    }


def vdp_param_ode():
    """VanDerPol with mu as an ODE parameter (PV = 1)."""
    args = oc.ODEArguments(2, 1, 1)
    x0, x1, u, mu = args.XVar(0), args.XVar(1), args.UVar(0), args.PVar(0)
    return oc.ODEBase(vf.stack([x1, (1.0 - x0 * x0) * x1 * mu - x0 + u]), 2, 1, 1)


def regions_phase(mode="LGL5", control_mode=None, nseg=6, accumulate=True):
    rng = np.random.default_rng(7)
    ode = vdp_param_ode()
    S = (W.CARDINALS[mode] - 1) * nseg + 1
    t = np.linspace(0.0, 2.0, S)
    T = np.stack([np.cos(t), -np.sin(t), t, 0.1 * np.sin(2 * t), np.full(S, 0.8)], axis=1)
    T[:, [0, 1, 3]] += 0.01 * rng.uniform(-1, 1, (S, 3))
    ph = ode.phase(mode)
    if control_mode:
        ph.setControlMode(control_mode)
    ph.setStaticParams([0.3, 1.7])
    ph.setTraj(T, nseg)
    A = vf.Arguments
    ph.addBoundaryValue("Front", [0, 1, 2], T[0, :3])
    ph.addEqualCon("Back", A(3).head2().squared_norm() - A(3)[2] * 0.1, [0, 1], [0], [])
    ph.addEqualCon("FrontandBack", A(4)[0] * A(4)[2] - A(4)[1] * A(4)[3], [0, 1])
    ph.addEqualCon("BackandFront", A(2)[0] - A(2)[1], [2])
    ph.addInequalCon("Path", vf.sin(A(3)[0]) * A(3)[1] - A(3)[2], [0, 3], [], [1])
    ph.addInequalCon("InnerPath", A(2).squared_norm() - 4.0, [0, 1])
    ph.addInequalCon("NodalPath", A(2)[0] * A(2)[1] - 1.0, [1, 3])
    ph.addInequalCon("PairWisePath", (A(4).head2() - A(4).tail2()).squared_norm() - 9.0, [0, 1])
    ph.addEqualCon("StaticParams", A(2)[0] * A(2)[1] - 0.51, [], [], [0, 1])
    ph.addEqualCon("ODEParams", vf.exp(A(1)[0]) - 2.2, [], [0], [])
    ph.addLUVarBound("Path", 3, -1.0, 1.0, 2.0)
    ph.addLUNormBound("Path", [0, 1], 0.01, 5.0)
    ph.addIntegralObjective(A(2).squared_norm() * A(2)[0], [1, 3])
    if accumulate:
        ph.addIntegralParamFunction(A(1)[0] * A(1)[0], [3], 1)
    ph.addStateObjective("Back", A(2)[0] * A(2)[1], [0, 2])
    ph.addDeltaTimeObjective(0.5)
    ph.addValueObjective("Front", 1, -1.0)
    return ph


def phase_zoo():
    """name -> builder returning a phase; small meshes, every region flag / transcription / control mode."""
    zoo = {}
    for mode in ("LGL3", "LGL5", "LGL7", "Trapezoidal"):
        zoo[f"lowthrust_{mode}"] = (lambda mode=mode: W.make_phase("lowthrust", mode, 7))
        zoo[f"vanderpol_{mode}"] = (lambda mode=mode: W.make_phase("vanderpol", mode, 9))
    zoo["hypersens_LGL5"] = lambda: W.make_phase("hypersens", "LGL5", 11)
    zoo["brachistochrone_LGL3"] = lambda: W.make_phase("brachistochrone", "LGL3", 8)
    zoo["lowthrust_LGL7_highest_order_spline"] = lambda: W.make_phase("lowthrust", "LGL7", 5, control_mode="HighestOrderSpline")
    zoo["lowthrust_LGL5_block_constant"] = lambda: W.make_phase("lowthrust", "LGL5", 6, control_mode="BlockConstant")
    zoo["lowthrust_LGL3_no_spline_defect_only"] = lambda: W.make_phase("lowthrust", "LGL3", 5, control_mode="NoSpline", extras=False)

    for mode in ("LGL3", "LGL5", "LGL7", "Trapezoidal"):
        zoo[f"regions_{mode}"] = (lambda mode=mode: regions_phase(mode))
    zoo["regions_LGL5_block_constant"] = lambda: regions_phase("LGL5", "BlockConstant")
    zoo["regions_LGL7_highest_order_spline"] = lambda: regions_phase("LGL7", "HighestOrderSpline")

    def ocp_multispacecraft(nphase=3, nseg=5):
        """Config 4 of BASELINE.json (examples/MultiSpacecraftOptimization.py:60-121): one BlockConstant LGL5 two-body
        phase per spacecraft, the back state and time of each tied to 7 free link parameters, plus a link-parameter
        constraint."""
        args = oc.ODEArguments(6, 3)
        r, v, u = args.head3(), args.segment3(3), args.tail3()
        ode = oc.ODEBase(vf.stack([v, r.normalized_power3() * (-1.0) + u * 0.01]), 6, 3)
        ocp = oc.OptimalControlProblem()
        rng = np.random.default_rng(11)
        for k in range(nphase):
            S = 2 * nseg + 1
            th = np.linspace(0.1 * k, 0.1 * k + 1.5, S)
            T = np.zeros((S, 10))
            T[:, 0], T[:, 1], T[:, 3], T[:, 4], T[:, 6] = np.cos(th), np.sin(th), -np.sin(th), np.cos(th), th
            T[:, 7:] = 0.3
            T[:, :6] += 0.01 * rng.uniform(-1, 1, (S, 6))
            ph = ode.phase("LGL5")
            ph.setControlMode("BlockConstant")
            ph.setTraj(T, nseg)
            ph.addBoundaryValue("Front", range(7), T[0, :7])
            ph.addLUNormBound("Path", [7, 8, 9], 0.01, 1.0)
            ph.addIntegralObjective(vf.Arguments(3).norm() * 0.1, [7, 8, 9])
            ocp.addPhase(ph)
        ocp.setLinkParams(np.array([0.1, 1.0, 0.02, -1.0, 0.1, 0.01, 1.6]))
        A = vf.Arguments
        link = A(14).head(7) - A(14).tail(7)
        for k in range(nphase):
            ocp.addLinkEqualCon(link, [(k, "Back", range(7), [], [])], range(7))
        ocp.addLinkParamEqualCon(A(6).head3().dot(A(6).tail3()), range(6))
        ocp.addLinkParamObjective(A(1)[0] * 0.5, [6])
        return ocp

    def ocp_forward_links():
        """Phases of different transcriptions chained by addForwardLinkEqualCon / addDirectLinkEqualCon, a flag-interface
        link over several phase pairs, a non-linear link inequality and a link objective."""
        ocp = oc.OptimalControlProblem()
        ode = vdp_param_ode()
        rng = np.random.default_rng(5)
        for k, (mode, nseg) in enumerate([("LGL3", 4), ("LGL5", 3), ("Trapezoidal", 5)]):
            S = (W.CARDINALS[mode] - 1) * nseg + 1
            t = np.linspace(k, k + 1.0, S)
            T = np.stack([np.cos(t), -np.sin(t), t, 0.1 * np.sin(2 * t), np.full(S, 0.8 + 0.1 * k)], axis=1)
            T[:, [0, 1, 3]] += 0.01 * rng.uniform(-1, 1, (S, 3))
            ph = ode.phase(mode)
            ph.setTraj(T, nseg)
            ph.addLUVarBound("Path", 3, -1.0, 1.0)
            ph.addIntegralObjective(vf.Arguments(1).squared_norm(), [3])
            ocp.addPhase(ph)
        A = vf.Arguments
        ocp.addForwardLinkEqualCon(0, -1, [0, 1, 2])
        ocp.addDirectLinkEqualCon(0, "ODEParams", [0], 1, "ODEParams", [0])
        ocp.addLinkEqualCon(A(2)[0] * A(2)[1] - 0.7, "BackToFront", [[0, 1], [1, 2]], [3])
        ocp.addLinkInequalCon(vf.sin(A(4)[0]) * A(4)[3] - A(4)[1] * A(4)[2], 0, "Front", [0, 1], 2, "Back", [0, 1])
        ocp.addLinkObjective(A(2)[1] - A(2)[0], 0, "Front", [2], 2, "Back", [2])
        return ocp

    zoo["ocp_multispacecraft_LGL5_block_constant"] = ocp_multispacecraft
    zoo["ocp_forward_links_mixed_transcriptions"] = ocp_forward_links
    return zoo


def problem_inputs(nlp, seed=3):
    d = nlp.dims()
    rng = np.random.default_rng(seed)
    LE = 100.0 * rng.uniform(-1, 1, d["neq"])
    LI = 100.0 * rng.uniform(-1, 1, d["niq"])
    sc = rng.uniform(0.5, 1.5, d["n_solver"])
    return LE, LI, sc
