"""The partial-matrix evaluations (evalKKTNO / evalSOE / evalAUG, NonLinearProgram.cpp:540-666) on the fused owner-writes
threads, opt-in through ASSET_B200_FUSED_PARTIAL=1.  Written when the round's GPU time was
spent: the kernels have been executed on the CPU emulator against the oracle (tests/test_ptx_emulation.py), the
plan-side flow (zero fill, solver-only entries, gathers, mesh ends) runs here for the first time -- hence the opt-in and
a file of its own, collected late."""
import numpy as np
import pytest

import oracle
from helpers import problem_inputs, rel_err, TOL
from asset_asrl_b200 import workloads as W

pytestmark = pytest.mark.gpu
MODES = {"KKT": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "KKTNO": ("AGX", "FXE", "FXI", "KKT"), "SOE": ("FXE", "FXI", "KKT"),
         "AUG": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "RHS": ("val", "PGX", "AGX", "FXE", "FXI"), "OCC": ("val", "FXE", "FXI")}
CASES = {
    "lowthrust_LGL5": lambda: W.make_phase("lowthrust", "LGL5", 400),
    "lowthrust_LGL7_highest_order_spline": lambda: W.make_phase("lowthrust", "LGL7", 300, control_mode="HighestOrderSpline"),
    "vanderpol_LGL3": lambda: W.make_phase("vanderpol", "LGL3", 500),
    "minclimb_LGL3": lambda: W.make_phase("minclimb", "LGL3", 300),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_partial_matrix_modes_on_the_fused_threads(name, monkeypatch):
    monkeypatch.setenv("ASSET_B200_FUSED_PARTIAL", "1")
    ph = CASES[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    calls = {"KKT": lambda: nlp.evalKKT(0.7, X, LE, LI), "KKTNO": lambda: nlp.evalKKTNO(0.7, X, LE, LI),
             "SOE": lambda: nlp.evalSOE(0.7, X, LE, LI), "AUG": lambda: nlp.evalAUG(0.7, X, LE, LI),
             "RHS": lambda: nlp.evalRHS(0.7, X, LE, LI), "OCC": lambda: nlp.evalOCC(0.7, X)}
    # interleaved, each mode at least three times (first call, graph capture, graph replay); a partial mode must leave no
    # stale value behind and must not disturb the next full evaluation
    for mode in ("KKT", "KKTNO", "KKT", "SOE", "AUG", "RHS", "KKTNO", "SOE", "AUG", "OCC", "KKTNO", "SOE", "AUG", "KKT"):
        out = calls[mode]()
        ref = orc.eval(mode, 0.7, X, LE, LI, sc)
        for k in MODES[mode]:
            if k == "val":
                assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (name, mode, k)
            else:
                assert rel_err(out[k], ref[k]) < TOL, (name, mode, k, rel_err(out[k], ref[k]))
        if mode in ("KKTNO", "SOE", "AUG"):
            assert nlp.fused_segments() > 0, (name, mode, "the fused kernels did not run")
    # new solver coefficients reach the entries that hold them alone, in a partial mode too
    sc2 = sc[::-1].copy()
    nlp.set_solver_coeffs(sc2)
    out = nlp.evalSOE(0.7, X, LE, LI)
    assert rel_err(out["KKT"], orc.eval("SOE", 0.7, X, LE, LI, sc2)["KKT"]) < TOL


def test_without_the_switch_the_partial_modes_stay_on_the_per_function_path(monkeypatch):
    monkeypatch.delenv("ASSET_B200_FUSED_PARTIAL", raising=False)
    ph = CASES["lowthrust_LGL5"]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    out = nlp.evalKKTNO(0.7, X, LE, LI)
    assert nlp.fused_segments() == 0
    ref = oracle.Oracle.from_problem(nlp).eval("KKTNO", 0.7, X, LE, LI, sc)
    assert rel_err(out["KKT"], ref["KKT"]) < TOL
