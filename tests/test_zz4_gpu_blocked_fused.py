"""Block-constant control on the fused owner-writes threads (ASSET_B200_FUSED_BLOCKED=1):
the per-segment controls are satellite variables of the threads (fused.cpp: Finder::satellites).  Written when the
round's GPU time was spent: the kernels and the whole data flow around them are checked on the CPU
(tests/test_fused_flow_model.py); this is their first run on a device -- opt-in, and a file collected last."""
import numpy as np
import pytest

import oracle
from helpers import problem_inputs, rel_err, TOL
from asset_asrl_b200 import workloads as W

pytestmark = pytest.mark.gpu
MODES = {"KKT": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "KKTNO": ("AGX", "FXE", "FXI", "KKT"), "SOE": ("FXE", "FXI", "KKT"),
         "AUG": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "RHS": ("val", "PGX", "AGX", "FXE", "FXI"),
         "OGC": ("val", "PGX", "FXE", "FXI"), "OCC": ("val", "FXE", "FXI"), "OBJ": ("val",)}


@pytest.mark.parametrize("ode,mode,nseg", [("lowthrust", "LGL5", 400), ("lowthrust", "LGL7", 300), ("lowthrust", "LGL3", 500),
                                           ("vanderpol", "Trapezoidal", 500), ("lowthrust", "LGL5", 20000)])
def test_block_constant_phase_on_the_fused_threads(ode, mode, nseg, monkeypatch):
    monkeypatch.setenv("ASSET_B200_FUSED_BLOCKED", "1")
    ph = W.make_phase(ode, mode, nseg, control_mode="BlockConstant")
    nlp = ph.transcribe()
    assert nlp.fused_info()["ok"], nlp.fused_info()["why"]
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    calls = {"KKT": lambda: nlp.evalKKT(0.7, X, LE, LI), "KKTNO": lambda: nlp.evalKKTNO(0.7, X, LE, LI),
             "SOE": lambda: nlp.evalSOE(0.7, X, LE, LI), "AUG": lambda: nlp.evalAUG(0.7, X, LE, LI),
             "RHS": lambda: nlp.evalRHS(0.7, X, LE, LI), "OGC": lambda: nlp.evalOGC(0.7, X),
             "OCC": lambda: nlp.evalOCC(0.7, X), "OBJ": lambda: nlp.evalOBJ(0.7, X)}
    for m in ("KKT", "RHS", "KKT", "SOE", "KKT", "OCC", "AUG", "KKTNO", "OGC", "OBJ", "KKT"):
        out = calls[m]()
        ref = orc.eval(m, 0.7, X, LE, LI, sc, threads=oracle.hardware_threads())
        for k in MODES[m]:
            if k == "val":
                assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (ode, mode, m, k)
            else:
                assert rel_err(out[k], ref[k]) < TOL, (ode, mode, m, k, rel_err(out[k], ref[k]))
        if m == "KKT":
            assert nlp.fused_segments() > 0, "the fused kernels did not run"


def test_without_the_switch_block_constant_stays_on_the_per_function_path(monkeypatch):
    monkeypatch.delenv("ASSET_B200_FUSED_BLOCKED", raising=False)
    ph = W.make_phase("lowthrust", "LGL5", 400, control_mode="BlockConstant")
    nlp = ph.transcribe()
    assert not nlp.fused_info()["ok"]


@pytest.mark.parametrize("mode,nseg", [("LGL5", 300), ("Trapezoidal", 400), ("LGL7", 200)])
def test_block_constant_with_parameters_and_every_region_flag(mode, nseg, monkeypatch):
    """ODE parameter, static parameters, functions over every region flag (helpers.regions_phase without its accumulation)."""
    from helpers import regions_phase
    monkeypatch.setenv("ASSET_B200_FUSED_BLOCKED", "1")
    ph = regions_phase(mode, "BlockConstant", nseg, accumulate=False)
    nlp = ph.transcribe()
    assert nlp.fused_info()["ok"], nlp.fused_info()["why"]
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    for m, call in (("KKT", lambda: nlp.evalKKT(0.7, X, LE, LI)), ("RHS", lambda: nlp.evalRHS(0.7, X, LE, LI)),
                    ("KKT", lambda: nlp.evalKKT(0.7, X, LE, LI))):
        out = call()
        ref = orc.eval(m, 0.7, X, LE, LI, sc)
        for k in MODES[m]:
            if k == "val":
                assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (mode, m, k)
            else:
                assert rel_err(out[k], ref[k]) < TOL, (mode, m, k, rel_err(out[k], ref[k]))
    assert nlp.fused_segments() > 0
