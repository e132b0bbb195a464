"""The owner-writes fused interior (csrc/fused.cpp) against the CPU oracle and against the per-function path.

The meshes are long enough (>= 32 periodic segments) for the fused kernel to be used; every evaluation mode is run
after the fused evalKKT to check that the two paths can alternate on one plan."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import oracle  # noqa: E402
from helpers import phase_zoo, problem_inputs, rel_err, TOL  # noqa: E402

from asset_asrl_b200 import workloads as W  # noqa: E402

pytestmark = pytest.mark.gpu
KEYS = ("PGX", "AGX", "FXE", "FXI", "KKT")
CASES = {
    "lowthrust_LGL5": lambda: W.make_phase("lowthrust", "LGL5", 150),
    "lowthrust_LGL7": lambda: W.make_phase("lowthrust", "LGL7", 100),
    "lowthrust_LGL3": lambda: W.make_phase("lowthrust", "LGL3", 150),
    "lowthrust_Trapezoidal": lambda: W.make_phase("lowthrust", "Trapezoidal", 150),
    "lowthrust_LGL7_highest_order_spline": lambda: W.make_phase("lowthrust", "LGL7", 90, control_mode="HighestOrderSpline"),
    "lowthrust_LGL5_defect_only": lambda: W.make_phase("lowthrust", "LGL5", 120, extras=False),
    "vanderpol_LGL5": lambda: W.make_phase("vanderpol", "LGL5", 200),
    "vanderpol_LGL3": lambda: W.make_phase("vanderpol", "LGL3", 200),
    "hypersens_LGL5": lambda: W.make_phase("hypersens", "LGL5", 200),
    "brachistochrone_LGL3": lambda: W.make_phase("brachistochrone", "LGL3", 100),
    "lowthrust_LGL5_block_constant": lambda: W.make_phase("lowthrust", "LGL5", 100, control_mode="BlockConstant"),
}


def check(out, ref, tag):
    for k in KEYS:
        assert rel_err(out[k], ref[k]) < TOL, (tag, k, rel_err(out[k], ref[k]))
    assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (tag, "val")


@pytest.mark.parametrize("name", sorted(CASES))
def test_fused_evalKKT_matches_oracle(name):
    ph = CASES[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info()
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    ref = orc.eval("KKT", 0.7, X, LE, LI, sc)
    out = nlp.evalKKT(0.7, X, LE, LI)
    check(out, ref, name)
    if info["ok"]:
        assert nlp.fused_segments() == info["t_hi"] - info["t_lo"] > 0
    else:
        assert nlp.fused_segments() == 0 and info["why"]
    # alternate with the per-function modes on the same plan, then evalKKT again with another iterate
    out = nlp.evalRHS(0.7, X, LE, LI)
    r2 = orc.eval("RHS", 0.7, X, LE, LI, sc)
    for k in ("PGX", "AGX", "FXE", "FXI"):
        assert rel_err(out[k], r2[k]) < TOL, (name, "RHS", k)
    out = nlp.evalSOE(0.7, X, LE, LI)
    r3 = orc.eval("SOE", 0.7, X, LE, LI, sc)
    assert rel_err(out["KKT"], r3["KKT"]) < TOL
    rng = np.random.default_rng(5)
    X2 = X + 1e-3 * rng.uniform(-1, 1, X.size)
    out = nlp.evalKKT(1.3, X2, LE * 0.5, LI * 2.0)
    check(out, orc.eval("KKT", 1.3, X2, LE * 0.5, LI * 2.0, sc), name + ":second")


def test_fused_is_used_for_the_bench_workload_and_agrees_with_the_per_function_path():
    ph = W.make_phase("lowthrust", "LGL5", 400)
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    a = nlp.evalKKT(0.7, X, LE, LI)
    assert nlp.fused_segments() >= 380
    os.environ["ASSET_B200_FUSED"] = "0"
    try:
        nlp2 = ph.transcribe()
        nlp2.to_device(0)
    finally:
        del os.environ["ASSET_B200_FUSED"]
    nlp2.set_solver_coeffs(sc)
    b = nlp2.evalKKT(0.7, X, LE, LI)
    assert nlp2.fused_segments() == 0
    for k in KEYS:
        assert rel_err(a[k], b[k]) < TOL, k


@pytest.mark.parametrize("name", ["lowthrust_LGL5", "lowthrust_LGL7", "vanderpol_LGL3", "hypersens_LGL5", "brachistochrone_LGL3"])
def test_fused_residual_only_modes(name):
    """evalRHS / OGC / OCC / OBJ on the fused kernels (the evaluations of PSIOPT's line search), interleaved with evalKKT
    and repeated so that the graph replay of every mode is exercised; val included."""
    ph = CASES[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    calls = {"RHS": (lambda x: nlp.evalRHS(0.7, x, LE, LI), ("PGX", "AGX", "FXE", "FXI", "val")),
             "OGC": (lambda x: nlp.evalOGC(0.7, x), ("PGX", "FXE", "FXI", "val")),
             "OCC": (lambda x: nlp.evalOCC(0.7, x), ("FXE", "FXI", "val")),
             "OBJ": (lambda x: nlp.evalOBJ(0.7, x), ("val",)),
             "KKT": (lambda x: nlp.evalKKT(0.7, x, LE, LI), KEYS + ("val",))}
    rng = np.random.default_rng(11)
    for rep in range(3):
        Xr = X + 1e-3 * rep * rng.uniform(-1, 1, X.size)
        for mode in ("RHS", "KKT", "OCC", "OGC", "OBJ", "RHS"):
            fn, keys = calls[mode]
            out = fn(Xr)
            ref = orc.eval(mode, 0.7, Xr, LE, LI, sc)
            for k in keys:
                if k == "val":
                    assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (name, mode, rep, "val")
                else:
                    assert rel_err(out[k], ref[k]) < TOL, (name, mode, rep, k, rel_err(out[k], ref[k]))
            if mode != "KKT" and nlp.fused_info()["ok"]:
                assert nlp.fused_segments() > 0, (name, mode, "the residual-only mode did not run on the fused kernels")
