"""GPU parity of the interpolation-table ops (InterpTable1D .. 4D), the matrix functions and of the
minimum-time-to-climb workload whose ODE reads them, against the CPU oracle through the C ABI.

A file of its own, collected after the other GPU tests: these kernels (table ops in the interpreter, in the specialised
per-function kernels and in the fused kernels, with the table as module-scope data) were added when no GPU time was
left in the round, so they have been checked on the CPU only (program interpreter, PTX emulator, in-process assembly
for sm_100a); a failure here must not hide the tests before it.
"""
import numpy as np
import pytest

import oracle
from helpers import interp_zoo, matrix_zoo, table_phase_zoo, problem_inputs, rel_err, TOL
from asset_asrl_b200 import workloads as W
from asset_asrl_b200.Solvers import NonLinearProgram

pytestmark = pytest.mark.gpu
MODES = {"KKT": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "KKTNO": ("AGX", "FXE", "FXI", "KKT"), "SOE": ("FXE", "FXI", "KKT"),
         "AUG": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "RHS": ("val", "PGX", "AGX", "FXE", "FXI"),
         "OGC": ("val", "PGX", "FXE", "FXI"), "OCC": ("val", "FXE", "FXI"), "OBJ": ("val",)}


def check(out, ref, keys, tag):
    for k in keys:
        if k == "val":
            assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (tag, k)
        else:
            assert rel_err(out[k], ref[k]) < TOL, (tag, k, rel_err(out[k], ref[k]))


def _cases():
    # every entry through the specialised kernels; a cross-section also through the interpreter and over a short range
    also = ("interp_1d_uneven_scalar_cubic", "interp_1d_even_vector_linear", "interp_2d_mixed_cubic", "interp_3d_mixed_cubic",
            "interp_4d_mixed_linear", "mat_products", "mat_inverse_4_pivoting")
    out = []
    for z in interp_zoo() + matrix_zoo():
        out.append(pytest.param(*z, "specialised_kernels", 4099, id=z[0] + "-specialised_bulk"))
        if z[0] in also:
            out.append(pytest.param(*z, "interpreter", 37, id=z[0] + "-interpreter"))
        if z[0] in also[:4]:
            out.append(pytest.param(*z, "specialised_kernels", 37, id=z[0] + "-specialised_short_range"))
    return out


@pytest.mark.parametrize("name,fun,gen,engine,nappl", _cases())
def test_table_ops_on_device(name, fun, gen, engine, nappl, monkeypatch):
    """Every table flavour (1-D .. 4-D, cubic / linear, even / uneven axes, scalar / vector data) and the matrix functions
    (products, inverses with select-based pivoting) as an equality constraint
    through each execution engine: the program interpreter (ASSET_B200_JIT=0: table pool in global memory), the
    specialised kernels over a short range and over 4099 applications (bulk kernel + indexed ends; table as module data)."""
    if engine == "interpreter":
        monkeypatch.setenv("ASSET_B200_JIT", "0")
    rng = np.random.default_rng(23)
    ir, orr = fun.IRows(), fun.ORows()
    nlp = NonLinearProgram(ir * nappl, orr * nappl, 0)
    t = nlp.add_tape(fun, True)
    V = np.arange(ir * nappl, dtype=np.int32).reshape(nappl, ir)
    Cc = np.arange(orr * nappl, dtype=np.int32).reshape(nappl, orr)
    nlp.add_function(1, t, False, V, Cc)
    nlp.analyze()
    X = np.concatenate([gen(rng) for _ in range(nappl)])
    LE = rng.uniform(-1, 1, orr * nappl)
    nlp.to_device(0)
    sc = np.zeros(nlp.dims()["n_solver"])
    nlp.set_solver_coeffs(sc)
    out = nlp.evalKKT(1.0, X, LE, np.zeros(0))
    ref = oracle.Oracle.from_problem(nlp).eval("KKT", 1.0, X, LE, np.zeros(0), sc)
    check(out, ref, ("AGX", "FXE", "KKT"), name)
    assert (nlp.specialised_launches() == 0) == (engine == "interpreter")


@pytest.mark.parametrize("name", sorted(table_phase_zoo()))
def test_table_phases_match_oracle(name):
    ph = table_phase_zoo()[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    check(nlp.evalKKT(0.7, X, LE, LI), orc.eval("KKT", 0.7, X, LE, LI, sc), MODES["KKT"], name)


def test_minimum_time_to_climb_at_1000_segments():
    """the climb ODE (five 1-D tables, one 2-D table) over 10^3 LGL5 segments, all eight
    evaluation modes (matrix, residual-only and value-only ones), fused kernels included."""
    ph = W.make_phase("minclimb", "LGL5", 1000, control_mode="HighestOrderSpline")
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    calls = {"KKT": lambda: nlp.evalKKT(0.7, X, LE, LI), "KKTNO": lambda: nlp.evalKKTNO(0.7, X, LE, LI),
             "SOE": lambda: nlp.evalSOE(0.7, X, LE, LI), "AUG": lambda: nlp.evalAUG(0.7, X, LE, LI),
             "RHS": lambda: nlp.evalRHS(0.7, X, LE, LI), "OGC": lambda: nlp.evalOGC(0.7, X),
             "OCC": lambda: nlp.evalOCC(0.7, X), "OBJ": lambda: nlp.evalOBJ(0.7, X)}
    for mode in ("KKT", "RHS", "SOE", "OCC", "OBJ", "KKT"):  # (every mode's kernels are assembled once; six keep the run short)
        check(calls[mode](), orc.eval(mode, 0.7, X, LE, LI, sc), MODES[mode], f"minclimb:{mode}")
    assert nlp.fused_segments() > 0
