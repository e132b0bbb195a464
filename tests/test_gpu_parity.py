"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on identical iterates.

Tolerance: 1e-12 relative to the output block's max-norm (BASELINE.json north_star: "within 1e-12 relative (fp64) for
constraint, Jacobian and Hessian values"); sparsity pattern and index maps bit-exact.
"""
import json
import os

import numpy as np
import pytest

import oracle
from helpers import phase_zoo, problem_inputs, rel_err, TOL, expression_zoo
from asset_asrl_b200 import workloads as W
from asset_asrl_b200.Solvers import NonLinearProgram

pytestmark = pytest.mark.gpu
ZOO = phase_zoo()
MODES = {"KKT": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "KKTNO": ("AGX", "FXE", "FXI", "KKT"), "SOE": ("FXE", "FXI", "KKT"),
         "AUG": ("val", "PGX", "AGX", "FXE", "FXI", "KKT"), "RHS": ("val", "PGX", "AGX", "FXE", "FXI"),
         "OGC": ("val", "PGX", "FXE", "FXI"), "OCC": ("val", "FXE", "FXI"), "OBJ": ("val",)}


def check(out, ref, keys, tag):
    for k in keys:
        if k == "val":
            assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0), (tag, k)
        else:
            assert rel_err(out[k], ref[k]) < TOL, (tag, k, rel_err(out[k], ref[k]))


@pytest.mark.parametrize("name", sorted(ZOO), ids=sorted(ZOO))
def test_evalKKT_matches_oracle(name):
    ph = ZOO[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    orc = oracle.Oracle.from_problem(nlp)
    ip, ix, loc = orc.pattern()
    ip2, ix2 = nlp.pattern()
    assert np.array_equal(ip, ip2) and np.array_equal(ix, ix2) and np.array_equal(loc, nlp.kkt_locations())
    out = nlp.evalKKT(0.7, X, LE, LI)
    check(out, orc.eval("KKT", 0.7, X, LE, LI, sc), MODES["KKT"], name)
    assert nlp.launches() > 0


@pytest.mark.parametrize("name", ["lowthrust_LGL5", "regions_LGL7", "regions_LGL5_block_constant", "vanderpol_Trapezoidal"])
def test_every_eval_mode(name):
    """The eight NonLinearProgram::eval* entry points (NonLinearProgram.cpp:347-683)."""
    ph = ZOO[name]()
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
    # interleave full and partial modes: a partial mode must not leave stale values behind
    for mode in ("KKT", "SOE", "KKT", "RHS", "AUG", "KKTNO", "OGC", "OCC", "OBJ", "KKT"):
        ref = orc.eval(mode, 0.7, X, LE, LI, sc)
        check(calls[mode](), ref, MODES[mode], f"{name}:{mode}")


@pytest.mark.parametrize("name,fun,gen", expression_zoo(), ids=[z[0] for z in expression_zoo()])
def test_expression_ops_on_device(name, fun, gen):
    """Every tape opcode through the device interpreter: the function as an equality constraint over 37 applications
    (ragged: not a multiple of the applications-per-warp packing)."""
    rng = np.random.default_rng(17)
    ir, orr, nappl = fun.IRows(), fun.ORows(), 37
    nlp = NonLinearProgram(ir * nappl, orr * nappl, 0)
    t = nlp.add_tape(fun, True)
    V = np.arange(ir * nappl, dtype=np.int32).reshape(nappl, ir)
    Cc = np.arange(orr * nappl, dtype=np.int32).reshape(nappl, orr)
    nlp.add_function(1, t, False, V, Cc)
    nlp.analyze()
    X = np.concatenate([gen(rng) for _ in range(nappl)])
    LE = rng.uniform(-1, 1, orr * nappl)
    nlp.to_device(0)
    nlp.set_solver_coeffs(np.zeros(nlp.dims()["n_solver"]))
    out = nlp.evalKKT(1.0, X, LE, np.zeros(0))
    ref = oracle.Oracle.from_problem(nlp).eval("KKT", 1.0, X, LE, np.zeros(0), np.zeros(nlp.dims()["n_solver"]))
    check(out, ref, ("AGX", "FXE", "KKT"), name)


def test_golden_defects_on_device():
    """One-segment phases evaluated on the GPU against the 50-digit mpmath fixtures (independent of the oracle)."""
    from helpers import vdp_param_ode
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "collocation_functions.json")))
    odes = {"lowthrust": W.lowthrust_ode, "brachistochrone": W.brachistochrone_ode, "vanderpol_mu": vdp_param_ode}
    for c in golden:
        if c["kind"] != "defect":
            continue
        ode = odes[c["ode"]]()
        S, n = c["CS"], c["XV"] + 1 + c["UV"]
        x = np.array(c["x"])
        T = np.zeros((S, n + c["PV"]))
        T[:, :n] = x[:S * n].reshape(S, n)
        if c["PV"]:
            T[:, n:] = x[S * n:]
        ph = ode.phase(c["mode"], T, 1)
        ph.setControlMode("NoSpline")
        nlp = ph.transcribe()
        d = nlp.dims()
        assert d["nfuncs"] <= 2  # defect (+ LGL mesh spacing for LGL5/7)
        lam = np.zeros(d["neq"])
        lam[:len(c["lam"])] = c["lam"]
        nlp.to_device(0)
        nlp.set_solver_coeffs(np.zeros(d["n_solver"]))
        out = nlp.evalKKT(1.0, ph.makeSolverInput(), lam, np.zeros(0))
        ir, orr = len(c["x"]), len(c["lam"])
        J, H = np.array(c["jx"]), np.array(c["hx"])
        ip, ix = nlp.pattern()
        dense = np.zeros((d["kktdim"], d["kktdim"]))
        for r in range(d["kktdim"]):
            dense[r, ix[ip[r]:ip[r + 1]]] = out["KKT"][ip[r]:ip[r + 1]]
        assert rel_err(out["FXE"][:orr], c["fx"]) < TOL, c["name"]
        assert rel_err(dense[:ir, d["nvars"]:d["nvars"] + orr].T, J) < TOL, c["name"]   # Jacobian block (upper triangle: J^T)
        assert rel_err(np.triu(dense[:ir, :ir]), np.triu(H)) < TOL, c["name"]


def test_full_size_properties():
    """Size-independent properties at a BASELINE-scale mesh (2*10^5 low-thrust LGL5 segments):
    determinism run to run (the fixed-order gather replaces the reference's mutex scatter), linearity of AGX and of the
    KKT Hessian part in the multipliers, FX independent of the multipliers, and a sampled comparison of segment blocks
    with the oracle evaluated on a window of the same mesh."""
    nseg = 200000
    ph = W.make_phase("lowthrust", "LGL5", nseg)
    nlp = ph.transcribe()
    d = nlp.dims()
    X = ph.makeSolverInput()
    LE, LI = W.multipliers(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(np.zeros(d["n_solver"]))
    a = nlp.evalKKT(1.0, X, LE, LI)
    b = nlp.evalKKT(1.0, X, LE, LI)
    for k in ("KKT", "FXE", "FXI", "PGX", "AGX"):
        assert np.array_equal(a[k], b[k]), k  # bit-exact determinism
    assert a["val"] == b["val"]
    c = nlp.evalKKT(1.0, X, 2.0 * LE, 2.0 * LI)
    z = nlp.evalKKT(1.0, X, 0.0 * LE, 0.0 * LI)
    assert np.array_equal(c["FXE"], a["FXE"]) and np.array_equal(c["FXI"], a["FXI"])
    assert rel_err(c["AGX"], 2.0 * a["AGX"]) < 1e-13
    assert rel_err(c["KKT"] - z["KKT"], 2.0 * (a["KKT"] - z["KKT"])) < 1e-12
    assert np.all(np.isfinite(a["KKT"]))
    # a window of 50 segments from the middle of the mesh as its own phase: defect rows must agree with the big mesh
    w0, nw = nseg // 2, 50
    S0 = 2 * w0
    traj = np.array(ph.returnTraj())[S0:S0 + 2 * nw + 1]
    phw = W.lowthrust_ode().phase("LGL5", traj, nw)
    nlpw = phw.transcribe()
    orc = oracle.Oracle.from_problem(nlpw)
    LEw = np.zeros(nlpw.dims()["neq"])
    LEw[:12 * nw] = LE[12 * w0:12 * (w0 + nw)]
    refw = orc.eval("KKT", 1.0, phw.makeSolverInput(), LEw, np.zeros(nlpw.dims()["niq"]))
    assert rel_err(a["FXE"][12 * w0:12 * (w0 + nw)], refw["FXE"][:12 * nw]) < TOL


def test_full_size_every_output_against_the_oracle():
    """The bench workload at its benchmarked size (2*10^5 low-thrust LGL5 segments, fused kernels): every output array of
    evalKKT -- KKT values, FXE, FXI, PGX, AGX, val -- against the oracle evaluated on the whole problem, and the CSR
    pattern bit-exact."""
    nseg = 200000
    ph = W.make_phase("lowthrust", "LGL5", nseg)
    nlp = ph.transcribe()
    d = nlp.dims()
    X = ph.makeSolverInput()
    LE, LI = W.multipliers(nlp)
    sc = np.random.default_rng(5).uniform(0.5, 1.5, d["n_solver"])
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    out = nlp.evalKKT(0.7, X, LE, LI)
    assert nlp.fused_segments() > 0.99 * nseg
    orc = oracle.Oracle.from_problem(nlp)
    ip, ix, loc = orc.pattern()
    ip2, ix2 = nlp.pattern()
    assert np.array_equal(ip, ip2) and np.array_equal(ix, ix2) and np.array_equal(loc, nlp.kkt_locations())
    ref = orc.eval("KKT", 0.7, X, LE, LI, sc, threads=oracle.hardware_threads())
    check(out, ref, MODES["KKT"], "full size")
    # block-wise as well: a wrong value in a small block must not hide behind the largest entry of the whole array
    nb = 4096
    for k in ("KKT", "AGX", "FXE"):
        a, b = out[k], ref[k]
        n = a.size // nb * nb
        blk = np.abs(b[:n]).reshape(-1, nb).max(axis=1)
        err = np.abs(a[:n] - b[:n]).reshape(-1, nb).max(axis=1)
        assert np.all(err <= TOL * np.maximum(blk, 1e-300) + 1e-300), (k, float((err / np.maximum(blk, 1e-300)).max()))


def test_full_size_oracle_windows():
    """The same mesh through oracle windows (oracle/windows.py): both mesh ends, the middle, and the seven cuts of a
    world-8 partition of the fused segments, 256 segments each."""
    from oracle.windows import check_windows
    nseg = 200000
    ph = W.make_phase("lowthrust", "LGL5", nseg)
    nlp = ph.transcribe()
    d = nlp.dims()
    X = ph.makeSolverInput()
    LE, LI = W.multipliers(nlp)
    sc = np.random.default_rng(5).uniform(0.5, 1.5, d["n_solver"])
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    out = nlp.evalKKT(0.7, X, LE, LI)
    vs = 20  # variables per segment: two nodes of 10
    centres = [128, nseg // 2, nseg - 128] + [nseg * r // 8 for r in range(1, 8)]
    windows = [(max(c - 128, 0) * vs, min(c + 128, nseg) * vs) for c in centres]
    worst = check_windows(nlp, out, windows, 0.7, X, LE, LI, sc)
    for k in ("KKT", "AGX", "PGX", "FXE", "FXI"):
        assert worst[k] < TOL, (k, worst)
