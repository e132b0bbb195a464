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
from helpers import expression_zoo, rel_err, vf, oc, TOL
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


@pytest.mark.parametrize("name,fun,gen", expression_zoo(), ids=[z[0] for z in expression_zoo()])
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
