"""The bench workloads built for the oracle WITHOUT the product (test infrastructure; used by bench.py's reference arm and
by tests/test_standalone_oracle.py).  Nothing here imports asset_asrl_b200 or loads libasset_b200.so: the phase layout and
the index maps come from oracle/indexing.py (the independent restatement of PhaseIndexer / transcribe_phase), the
functions are written down as the s-expressions the oracle parses, the trajectory and multipliers are generated with the
same numpy recipe as asset_asrl_b200/workloads.py.  So the reference arm's problem set-up does not
run through the code it is compared with, and the test can ask whether two independent builds of the same workload
agree on the sparsity pattern, KKTLocations and every value.

Supported: the four example ODEs x LGL3 / LGL5 / LGL7 / Trapezoidal x the spline control modes, with or without the
example functions (bounds, norm bound, integral objective, boundary values).  Block-constant control is not (its solver
input is laid out by the product's PhaseIndexer code).
"""
import numpy as np

from . import indexing as RI
from .binding import Oracle

CARDINALS = {"LGL3": 2, "LGL5": 3, "LGL7": 4, "Trapezoidal": 2}
TMODE = {"LGL3": RI.LGL3, "LGL5": RI.LGL5, "LGL7": RI.LGL7, "Trapezoidal": RI.Trapezoidal}
CMODE = {None: RI.FirstOrderSpline, "FirstOrderSpline": RI.FirstOrderSpline, "HighestOrderSpline": RI.HighestOrderSpline,
         "NoSpline": RI.NoSpline}
SIZES = {"lowthrust": (6, 3), "vanderpol": (2, 1), "hypersens": (1, 1), "brachistochrone": (3, 1)}  # XV, UV


def S(kind, ir, orows, ip=(), d=(), *children):
    """One node of the oracle's s-expression format: (kind ir orows linear-flag [ints] [doubles] children...)."""
    return "(%s %d %d 0 [%s] [%s]%s)" % (kind, ir, orows, " ".join(str(int(i)) for i in ip), " ".join(repr(float(v)) for v in d),
                                         "".join(" " + c for c in children))


def _args(n):
    return S("args", n, n)


def _seg(n, start, size):
    return S("segment", n, size, (start, size), (), _args(n))


def ode_expression(name):
    """The example ODEs over [x, t, u] (examples/SimpleLowThrust.py:32-39, VanDerPol.py:20-27, HyperSens.py:23-26,
    Brachistochrone.py:23-31), same constants as asset_asrl_b200/workloads.py."""
    if name == "lowthrust":  # [v ; -r/|r|^3 + 0.02 u]
        n = 10
        grav = S("scaled", n, 3, (), (-1.0,), S("normalized", n, 3, (3,), (), _seg(n, 0, 3)))
        return S("stack", n, 6, (), (), _seg(n, 3, 3), S("sum", n, 3, (), (), grav, S("scaled", n, 3, (), (0.02,), _seg(n, 7, 3))))
    if name == "vanderpol":  # [x1 ; (1 - x0^2) x1 - x0 + u]
        n = 4
        x0, x1, u = _seg(n, 0, 1), _seg(n, 1, 1), _seg(n, 3, 1)
        damp = S("cwiseprod", n, 1, (), (), S("vecminus", n, 1, (), (1.0,), S("cwiseprod", n, 1, (), (), x0, x0)), x1)
        return S("stack", n, 2, (), (), x1, S("sum", n, 1, (), (), S("diff", n, 1, (), (), damp, x0), u))
    if name == "hypersens":  # [u - x]
        n = 3
        return S("stack", n, 1, (), (), S("diff", n, 1, (), (), _seg(n, 2, 1), _seg(n, 0, 1)))
    if name == "brachistochrone":  # [v sin(theta) ; -v cos(theta) ; g cos(theta)]; unary 16 = sin, 17 = cos (csrc/tape.h)
        n = 5
        v, th = _seg(n, 2, 1), _seg(n, 4, 1)
        sin, cos = S("unary", n, 1, (16,), (), th), S("unary", n, 1, (17,), (), th)
        return S("stack", n, 3, (), (), S("cwiseprod", n, 1, (), (), sin, v), S("scaled", n, 1, (), (-1.0,), S("cwiseprod", n, 1, (), (), cos, v)),
                 S("scaled", n, 1, (), (9.81,), cos))
    raise ValueError(f"unknown workload ODE '{name}'")


def reference_trajectory(name, npts, rng):
    """Node states [x, t, u] of the measurement workloads: asset_asrl_b200/workloads.py's recipe (smooth orbit / oscillation
    perturbed by 1e-2 U(-1, 1), seed 12345)."""
    def perturb(T, cols, scale):
        T[:, cols] += scale * rng.uniform(-1.0, 1.0, (T.shape[0], len(cols)))
    if name == "lowthrust":
        th = np.linspace(0.0, 6.4 * np.pi, npts)
        T = np.zeros((npts, 10))
        T[:, 0], T[:, 1] = np.cos(th), np.sin(th)
        T[:, 3], T[:, 4] = -np.sin(th), np.cos(th)
        T[:, 6] = th
        T[:, 7], T[:, 8] = -0.8 * np.sin(th), 0.8 * np.cos(th)
        perturb(T, list(range(6)) + [7, 8, 9], 1e-2)
    elif name == "vanderpol":
        t = np.linspace(0.0, 10.0, npts)
        T = np.stack([np.cos(t), -np.sin(t), t, 0.1 * np.sin(2 * t)], axis=1)
        perturb(T, [0, 1, 3], 1e-2)
    elif name == "hypersens":
        t = np.linspace(0.0, 1.0e4, npts)
        T = np.stack([1.0 + 0.5 * np.sin(t * 1e-3), t, 0.5 * np.cos(t * 1e-3)], axis=1)
        perturb(T, [0, 2], 1e-2)
    elif name == "brachistochrone":
        t = np.linspace(0.0, 1.0, npts)
        T = np.stack([t, 1.0 - t, 1.0 + 3.0 * t, t, 0.3 + t], axis=1)
        perturb(T, [0, 1, 2, 4], 1e-2)
    else:
        raise ValueError(f"unknown workload ODE '{name}'")
    return T


class _Fun:
    def __init__(self, desc, orows, linear):
        self.desc, self.orows, self.linear = desc, orows, linear


class _Ode:
    def __init__(self, XV, UV):
        self.XV, self.UV, self.PV = XV, UV, 0


class _Phase:
    """What oracle/indexing.py reads of a phase: sizes, modes, the functions as the user added them."""

    def __init__(self, XV, UV, tmode, cmode, nseg):
        self.ode, self._spv, self.tmode, self._cmode, self._imode, self._nseg = _Ode(XV, UV), 0, tmode, cmode, RI.BaseIntegral, nseg
        self._funcs = []

    def add(self, role, fun, reg, x):
        self._funcs.append((role, fun, reg, list(x), [], [], []))


def _lu_bound(lb, ub):
    """StackedOutputs{(lb - x), (x - ub)} (ODEPhaseBase.cpp:110-130), unit scales."""
    x = _args(1)
    return S("stack", 1, 2, (), (), S("scaled", 1, 1, (), (1.0,), S("vecminus", 1, 1, (), (lb,), x)),
             S("scaled", 1, 1, (), (1.0,), S("addvec", 1, 1, (), (-ub,), x)))


def _boundary(values):
    n = len(values)
    return _Fun("expr " + S("addvec", n, n, (), [-v for v in values], _args(n)), n, True)


def build(name, mode, nseg, control_mode=None, extras=True, seed=12345):
    """(oracle problem, X, {"nvars", "neq", "niq"}) of asset_asrl_b200.workloads.make_phase(name, mode, nseg, ...)."""
    funcs, X, dims = describe(name, mode, nseg, control_mode, extras, seed)
    return instantiate(funcs, dims), X, dims


def instantiate(funcs, dims):
    """An oracle problem (of the currently selected build, oracle.use_variant) from describe()'s function list."""
    orc = Oracle(dims["nvars"], dims["neq"], dims["niq"])
    for role, desc, linear, V, Cc, appconst in funcs:
        orc.add_function(role, desc, linear, V, Cc, appconst)
    orc.analyze()
    return orc


def describe(name, mode, nseg, control_mode=None, extras=True, seed=12345):
    """([(role, description, linear, Vindex, Cindex, per-application constants)], X, dims): the workload as data."""
    if control_mode == "BlockConstant":
        raise NotImplementedError("block-constant control is not covered by the standalone builder")
    XV, UV = SIZES[name]
    CS = CARDINALS[mode]
    tmode, cmode = TMODE[mode], CMODE[control_mode]
    rng = np.random.default_rng(seed)
    npts = (CS - 1) * nseg + 1
    T = reference_trajectory(name, npts, rng)
    ph = _Phase(XV, UV, tmode, cmode, nseg)
    if extras:
        ph.add(1, _boundary(T[0, :XV + 1]), RI.Front, range(XV + 1))
        ph.add(1, _boundary(T[-1, :XV]), RI.Back, range(XV))
        ucols = list(range(XV + 1, XV + 1 + UV))
        if name == "lowthrust":
            nb = S("nested", 3, 2, (), (), _lu_bound(0.001, 1.0), S("norm", 3, 1, (1,), (), _args(3)))
            ph.add(2, _Fun("expr " + nb, 2, False), RI.Path, ucols)
            integrand, nu = S("scaled", 3, 1, (), (0.5,), S("norm", 3, 1, (2,), (), _args(3))), 3
            ph.add(3, _Fun(integrand, 1, False), RI.Path, ucols)
        else:
            ph.add(2, _Fun("expr " + _lu_bound(-10.0, 10.0), 2, True), RI.Path, ucols[:1])
            integrand, nu = S("scaled", 1, 1, (), (0.5,), S("norm", 1, 1, (2,), (), _args(1))), 1
            ph.add(3, _Fun(integrand, 1, False), RI.Path, ucols[:1])
    lists, dims = RI.transcribe_single_phase(ph)
    # the functions behind the names oracle/indexing.py emits, in the reference's emission order
    ode = ode_expression(name)
    defect = ("trap_defect" if mode == "Trapezoidal" else "lgl_defect") + f" {CS} {XV} {UV} 0 " + ode
    order = 2 if (tmode == RI.LGL7 and cmode == RI.HighestOrderSpline) else 1
    user = {1: iter([f for (r, f, *_rest) in ph._funcs if r == 1]), 2: iter([f for (r, f, *_rest) in ph._funcs if r == 2]),
            3: iter([f for (r, f, *_rest) in ph._funcs if r == 3])}
    funcs = []

    def add(role, desc, linear, V, Cc=None, appconst=None):
        funcs.append((role, desc, linear, np.asarray(V, dtype=np.int32), None if Cc is None else np.asarray(Cc, dtype=np.int32), appconst))

    for fn in lists.obj:  # objectives first, then equalities, then inequalities: the solver's evaluation order
        f = next(user[3])
        add(0, f"lgl_integral {CS} {nu} 0 " + f.desc, False, fn["Vindex"])
    singles = []

    def flush_singles():
        # the reference emits N-1 single-application SingleMeshSpacing(c_i) functions (addPartitionedEquality,
        # PhaseIndexer.cpp:20-39); one function with the constant per application evaluates the same triplets in the same order
        if not singles:
            return
        c = np.linspace(0.0, 1.0, nseg + 1)[1:nseg]  # ODEPhaseBase.cpp:963-969 (one bin)
        assert len(singles) == c.size
        add(1, "single_mesh_spacing", True, np.concatenate([f["Vindex"] for f in singles]), np.concatenate([f["Cindex"] for f in singles]),
            c.reshape(-1, 1).copy())
        singles.clear()

    for fn in lists.eq:
        nm = fn["name"]
        if nm == "single_mesh_spacing":
            singles.append(fn)
            continue
        flush_singles()
        if nm == "defect":
            add(1, defect, False, fn["Vindex"], fn["Cindex"])
        elif nm == "lgl_mesh_spacing":
            add(1, f"lgl_mesh_spacing {CS}", False, fn["Vindex"], fn["Cindex"])
        elif nm == "control_spline":
            add(1, f"control_spline {CS} {UV} {order}", False, fn["Vindex"], fn["Cindex"])
        elif nm == "user_equality":
            f = next(user[1])
            add(1, f.desc, f.linear, fn["Vindex"], fn["Cindex"])
        else:
            raise AssertionError("unexpected equality " + nm)
    flush_singles()
    for fn in lists.iq:
        f = next(user[2])
        add(2, f.desc, f.linear, fn["Vindex"], fn["Cindex"])
    X = np.ascontiguousarray(T).ravel().copy()  # non-blocked layout: node states back to back (PhaseIndexer.h:63-99), no parameters
    return funcs, X, dims


def multipliers(dims, seed=67890):
    """100 U(-1, 1) (the reference's NLPTest choice, NonLinearProgram.cpp:708-719), as asset_asrl_b200/workloads.py draws them."""
    rng = np.random.default_rng(seed)
    return 100.0 * rng.uniform(-1, 1, dims["neq"]), 100.0 * rng.uniform(-1, 1, dims["niq"])
