"""The specialiser's PTX executed on the CPU (tests/ptx_emu.py): the fused owner-writes kernels of several phases, a few
threads each, against the CPU oracle -- no GPU and no assembler involved, so an emitter bug (wrong register, wrong store
phase, a value parked before it is defined) shows up here, and an assembler bug does not.  The assembler's handling
of the parked-value accesses is checked separately on the SASS (tools/sass_park_order.py)."""
import os
import shutil
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))
import oracle  # noqa: E402
from emu_harness import emulate_fused  # noqa: E402
from helpers import problem_inputs  # noqa: E402

from asset_asrl_b200 import workloads as W  # noqa: E402

CASES = {
    "lowthrust_LGL5": lambda: W.make_phase("lowthrust", "LGL5", 60),
    "lowthrust_LGL7": lambda: W.make_phase("lowthrust", "LGL7", 50),
    "lowthrust_LGL7_highest_order_spline": lambda: W.make_phase("lowthrust", "LGL7", 50, control_mode="HighestOrderSpline"),
    "lowthrust_LGL3": lambda: W.make_phase("lowthrust", "LGL3", 60),
    "lowthrust_Trapezoidal": lambda: W.make_phase("lowthrust", "Trapezoidal", 60),
    "vanderpol_LGL5": lambda: W.make_phase("vanderpol", "LGL5", 60),
    "hypersens_LGL5": lambda: W.make_phase("hypersens", "LGL5", 60),
    "brachistochrone_LGL3": lambda: W.make_phase("brachistochrone", "LGL3", 60),
    # interpolation tables inside the ODE: table ops and their module-scope data in the fused kernels
    "minclimb_LGL5_highest_order_spline": lambda: W.make_phase("minclimb", "LGL5", 60, control_mode="HighestOrderSpline"),
    "minclimb_LGL3": lambda: W.make_phase("minclimb", "LGL3", 60),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_fused_ptx_on_the_emulator_matches_the_oracle(name):
    ph = CASES[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    ref = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc)
    R = info["t_hi"] - info["t_lo"]
    res = emulate_fused(nlp, X, LE, LI, sc, 0.7, [0, 1, 2, 3, R // 2, R - 1])  # every store phase (rel mod 4) and both ends
    wrote = 0
    res.pop("stage")
    for k, (v, m) in res.items():
        if not m.any():
            continue
        wrote += int(m.sum())
        scale = max(float(np.abs(ref[k]).max()), 1e-300)
        assert float(np.abs(v[m] - ref[k][m]).max()) <= 1e-12 * scale, (name, k)
    assert wrote >= 6 * info["slots_per_segment"]


@pytest.mark.parametrize("mode,name", [(4, "RHS"), (5, "OGC"), (6, "OCC"), (7, "OBJ")])
@pytest.mark.parametrize("case", ["lowthrust_LGL5", "vanderpol_LGL5", "lowthrust_LGL7_highest_order_spline", "minclimb_LGL3"])
def test_residual_only_fused_ptx_matches_the_oracle(case, mode, name):
    """The fused kernels of the line-search evaluations (evalRHS / OGC / OCC / OBJ): what a thread writes -- constraint
    values, gradient entries of its rows -- against the oracle in that mode; no matrix entry may be written."""
    ph = CASES[case]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info()
    ptx = nlp.fused_residual_ptx(mode)
    if ptx is None:  # nothing for the interior threads to do: the climb phase's objective touches the mesh ends only
        assert case.startswith("minclimb") and name == "OBJ"
        return
    assert ("ab_fused_m%d_0" % mode) in ptx
    assert nlp.assemble_ptx(ptx)["cubin_bytes"] > 0
    ref = oracle.Oracle.from_problem(nlp).eval(name, 0.7, X, LE, LI, sc)
    R = info["t_hi"] - info["t_lo"]
    res = emulate_fused(nlp, X, LE, LI, sc, 0.7, [0, 1, 2, 3, R // 2, R - 1], residual_mode=mode)
    res.pop("stage")
    assert not res["KKT"][1].any()
    expect = {"RHS": ("FXE", "FXI", "AGX", "PGX"), "OGC": ("FXE", "FXI", "PGX"), "OCC": ("FXE", "FXI"), "OBJ": ()}[name]
    for k, (v, m) in res.items():
        if k == "KKT":
            continue
        if k in ("FXE", "AGX") and k in expect:
            assert m.any(), (case, name, k, "the interior threads own constraint rows and gradient entries of this array")
        if m.any():
            assert k in expect, (case, name, k, "an array this mode does not produce was written")
            scale = max(float(np.abs(ref[k]).max()), 1e-300)
            assert float(np.abs(v[m] - ref[k][m]).max()) <= 1e-12 * scale, (case, name, k)


@pytest.mark.skipif(shutil.which("ptxas") is None or shutil.which("cuobjdump") is None, reason="needs the CUDA toolkit's ptxas and cuobjdump")
@pytest.mark.parametrize("name", ["lowthrust_LGL5", "lowthrust_LGL7_highest_order_spline"])
def test_assembler_keeps_the_order_of_parked_value_accesses(name):
    import sass_park_order
    nlp = CASES[name]().transcribe()
    info = nlp.fused_info()
    for p in range(int(info["pieces"])):
        assert sass_park_order.check(nlp.fused_ptx(p)) == [], (name, p)


def _single_function_cases():
    from helpers import interp_zoo, matrix_zoo
    want = ("interp_1d_uneven_scalar_cubic", "interp_2d_mixed_cubic", "interp_3d_mixed_cubic", "interp_4d_mixed_cubic", "interp_4d_mixed_linear",
            "mat_products", "mat_inverse_4_pivoting")
    return [z for z in interp_zoo() + matrix_zoo() if z[0] in want]


@pytest.mark.parametrize("name,fun,gen", _single_function_cases(), ids=[z[0] for z in _single_function_cases()])
def test_fused_ptx_of_table_and_matrix_functions_matches_the_oracle(name, fun, gen):
    """One function over many applications is a fusable problem too: its fused kernel (table ops, table data copied into the
    fused DAG; select-based pivoting of the matrix inverse) on the emulator against the oracle."""
    from asset_asrl_b200.Solvers import NonLinearProgram
    rng = np.random.default_rng(31)
    ir, orr, nappl = fun.IRows(), fun.ORows(), 200
    nlp = NonLinearProgram(ir * nappl, orr * nappl, 0)
    t = nlp.add_tape(fun, True)
    nlp.add_function(1, t, False, np.arange(ir * nappl, dtype=np.int32).reshape(nappl, ir), np.arange(orr * nappl, dtype=np.int32).reshape(nappl, orr))
    nlp.analyze()
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    X = np.concatenate([gen(rng) for _ in range(nappl)])
    LE = rng.uniform(-1, 1, orr * nappl)
    sc = np.linspace(0.5, 1.5, nlp.dims()["n_solver"])
    ref = oracle.Oracle.from_problem(nlp).eval("KKT", 1.0, X, LE, np.zeros(0), sc)
    R = info["t_hi"] - info["t_lo"]
    res = emulate_fused(nlp, X, LE, np.zeros(0), sc, 1.0, [0, 1, 2, 3, R // 2, R - 1])
    res.pop("stage")
    wrote = 0
    for k, (v, m) in res.items():
        if not m.any():
            continue
        wrote += int(m.sum())
        scale = max(float(np.abs(ref[k]).max()), 1e-300)
        assert float(np.abs(v[m] - ref[k][m]).max()) <= 1e-12 * scale, (name, k)
    assert wrote >= 6 * info["slots_per_segment"]


@pytest.mark.parametrize("mode,name", [(1, "KKTNO"), (2, "SOE"), (3, "AUG")])
@pytest.mark.parametrize("case", ["lowthrust_LGL5", "vanderpol_LGL5", "lowthrust_LGL3", "minclimb_LGL3"])
def test_partial_matrix_fused_ptx_matches_the_oracle(case, mode, name):
    """The fused kernels of evalKKTNO / evalSOE / evalAUG (NonLinearProgram.cpp:540-666): every owned matrix entry is written
    with the contributions the mode computes (zero where it computes none), the arrays the mode does not produce are
    left alone; values against the oracle evaluated in that mode."""
    if case == "minclimb_LGL3" and name != "KKTNO":
        pytest.skip("one mode of the table phase is enough for the CPU suite's time budget")
    ph = CASES[case]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info()
    ref = oracle.Oracle.from_problem(nlp).eval(name, 0.7, X, LE, LI, sc)
    R = info["t_hi"] - info["t_lo"]
    rels = [0, 1, 2, 3, R - 1]
    res = emulate_fused(nlp, X, LE, LI, sc, 0.7, rels, residual_mode=mode)
    res.pop("stage")
    p = 0
    while nlp.fused_mode_ptx(mode, p) is not None:
        assert nlp.assemble_ptx(nlp.fused_mode_ptx(mode, p))["cubin_bytes"] > 0
        p += 1
    assert res["KKT"][1].sum() >= len(rels) * info["slots_per_segment"]  # every owned matrix entry is written
    expect = {"KKTNO": ("KKT", "FXE", "FXI", "AGX"), "SOE": ("KKT", "FXE", "FXI"), "AUG": ("KKT", "FXE", "FXI", "AGX", "PGX")}[name]
    for k, (v, m) in res.items():
        if k not in expect:
            assert not m.any(), (case, name, k, "an array this mode does not produce was written")
            continue
        if k in ("KKT", "FXE") or (k == "AGX"):
            assert m.any(), (case, name, k)
        if m.any():
            scale = max(float(np.abs(ref[k]).max()), 1e-300)
            assert float(np.abs(v[m] - ref[k][m]).max()) <= 1e-12 * scale, (case, name, k)
