"""The fused evaluation's data flow, modelled on the CPU and checked against the oracle (no GPU).

The kernels alone are covered by tests/test_ptx_emulation.py; this test covers what csrc/plan.cu builds around them
for an evalKKT on one device (Plan::build's gather lists for the fused evaluation, Plan::enqueue):

  1. every fused thread runs (the emitted PTX on tests/ptx_emu.py): owned CSR entries, gradient entries and constraint
     values straight into the output arrays, entries of rows no thread owns either straight (one source) or into the
     staging buffer;
  2. the per-function programs (tests/prog_ref.py) evaluate the applications the fused threads do not
     (FusedPlan::ex_lo / ex_hi), writing through the destination maps -- except that a matrix entry a fused thread
     owns is not theirs to write (the redirected maps, Plan::kdF);
  3. the fixed-order gather lists sum the staging buffer into every target outside the owned rows.

The result must be the oracle's evalKKT.  Run for ordinary phases -- where the device agrees (tests/test_gpu_fused.py),
which validates the model -- and, with ASSET_B200_FUSED_BLOCKED=1, for block-constant control, whose per-segment
controls are satellite variables of the fused threads.
"""
import numpy as np
import pytest

import oracle
from emu_harness import emulate_fused
from helpers import problem_inputs
from ref_assembly import tape_block_eval
from asset_asrl_b200 import workloads as W


def model_fused_evalkkt(nlp, X, LE, LI, sc, obj_scale, mode=0):
    """mode 0: evalKKT; 1 / 2 / 3: evalKKTNO / evalSOE / evalAUG on the fused threads (ASSET_B200_FUSED_PARTIAL): the value
    array and the staging slots below the solver coefficients start from zero (Plan::enqueue), contributions the mode
    does not compute are never written."""
    with_obj = mode in (0, 3)          # objective functions are evaluated (gradient and value; Hessian only in evalKKT)
    with_hess = mode in (0, 1)
    with_grad = mode in (0, 1, 3)
    d = nlp.dims()
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    lay = nlp.fused_layout()
    R = info["t_hi"] - info["t_lo"]
    k_lo, k_hi = lay["kb"], lay["kb"] + R * lay["ks"]
    v_lo, v_hi = lay["vb"], lay["vb"] + R * lay["vs"]
    # 1. the fused threads
    res = emulate_fused(nlp, X, LE, LI, sc, obj_scale, range(R), residual_mode=mode or None)
    stage = res.pop("stage")
    out = {k: v.copy() for k, (v, _) in res.items()}
    direct_by_fused = {k: m.copy() for k, (_, m) in res.items()}
    assert direct_by_fused["KKT"][k_lo:k_hi].all(), "a fused thread leaves one of its own matrix entries unwritten"
    # 2. what is left to the per-function path
    ev = tape_block_eval(nlp)
    for f in range(d["nfuncs"]):
        fi = nlp.function_info(f)
        V, Cc = nlp.function_indices(f)
        kd, fd, g_stage, v_stage = nlp.dest_maps(f)
        appc_all = nlp.function_appconst(f)
        L = LE if fi["role"] == 1 else LI
        if fi["role"] == 0 and not with_obj:
            continue
        for a in range(fi["nappl"]):
            if lay["ex_lo"][f] <= a < lay["ex_hi"][f]:
                continue
            lam = np.array([obj_scale]) if fi["role"] == 0 else L[Cc[a]]
            fx, J, H = ev(f, fi, X[V[a]], lam, appc_all[a] if fi["nappconst"] else ())
            e = 0
            for i in range(fi["ir"]):
                hess = [] if fi["linear"] else [(H[j, i], with_hess and (fi["role"] != 0 or mode == 0)) for j in range(i, fi["ir"])]
                jac = [(J[k, i], True) for k in range(fi["orows"])] if fi["role"] else []
                for v, computed in hess + jac:
                    dst = int(kd[a, e])
                    e += 1
                    if not computed:
                        continue
                    if dst < 0:
                        stage[~dst] = v
                    elif not (k_lo <= dst < k_hi):  # a fused thread's own entry is redirected away (Plan::kdF)
                        if direct_by_fused["KKT"][dst]:
                            # the last fused thread's applications reach into rows no fused thread owns, so the per-function
                            # path evaluates them too; entries of far rows are then written by both, with the same value
                            assert abs(out["KKT"][dst] - v) <= 1e-13 * max(abs(v), 1.0), "two writers of a matrix entry disagree"
                        out["KKT"][dst] = v
            g = J.T @ lam
            for i in range(fi["ir"]):
                if with_grad:
                    stage[g_stage + a * fi["ir"] + i] = g[i]
            if fi["role"] == 0:
                stage[v_stage + a] = fx[0] * obj_scale
            else:
                arr = out["FXE" if fi["role"] == 1 else "FXI"]
                for k in range(fi["orows"]):
                    dst = int(fd[a, k])
                    if dst < 0:
                        stage[~dst] = fx[k]
                    else:
                        arr[dst] = fx[k]
    # 3. the gathers: everything outside the owned rows that has a list entry
    lists, _ = nlp.gather_lists()
    val = 0.0
    produced = {0: ("KKT", "PGX", "AGX", "FXE", "FXI", "val"), 1: ("KKT", "AGX", "FXE", "FXI"), 2: ("KKT", "FXE", "FXI"),
                3: ("KKT", "PGX", "AGX", "FXE", "FXI", "val")}[mode]
    for name, owned in (("KKT", (k_lo, k_hi)), ("PGX", (v_lo, v_hi)), ("AGX", (v_lo, v_hi)), ("FXE", None), ("FXI", None), ("val", None)):
        if name not in produced:
            continue
        for ptr, src, tgt in lists[name]:
            for t in range(tgt.size):
                if owned is not None and owned[0] <= tgt[t] < owned[1]:
                    continue
                s_ = float(stage[src[ptr[t]:ptr[t + 1]]].sum())
                if name == "val":
                    val = s_
                else:
                    out[name][tgt[t]] = s_
    # what nobody writes keeps the zero the plan filled the arrays with when it was built: structurally zero outputs are
    # not stored (jit.cpp: skip_zero_outputs); a missing non-zero shows up in the comparison with the oracle
    sizes = {"KKT": d["nnz"], "PGX": d["nvars"], "AGX": d["nvars"], "FXE": d["neq"], "FXI": d["niq"]}
    for k in sizes:
        out[k] = np.nan_to_num(out[k], nan=0.0)[:sizes[k]]
    out["val"] = val
    return out


def check(nlp, ph, tag, mode=0):
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    got = model_fused_evalkkt(nlp, X, LE, LI, sc, 0.7, mode)
    name = ("KKT", "KKTNO", "SOE", "AUG")[mode]
    ref = oracle.Oracle.from_problem(nlp).eval(name, 0.7, X, LE, LI, sc)
    arrays = {0: ("KKT", "FXE", "FXI", "PGX", "AGX"), 1: ("KKT", "FXE", "FXI", "AGX"), 2: ("KKT", "FXE", "FXI"), 3: ("KKT", "FXE", "FXI", "PGX", "AGX")}[mode]
    for k in arrays:
        if ref[k].size:
            scale = max(float(np.abs(ref[k]).max()), 1e-300)
            assert float(np.abs(got[k] - ref[k]).max()) <= 1e-12 * scale, (tag, name, k)
    if mode in (0, 3):
        assert abs(got["val"] - ref["val"]) <= 1e-12 * max(abs(ref["val"]), 1.0), (tag, name)


@pytest.mark.parametrize("ode,mode,nseg,control", [("lowthrust", "LGL5", 44, None), ("vanderpol", "LGL3", 48, None),
                                                   ("lowthrust", "LGL7", 40, "HighestOrderSpline")])
def test_model_of_the_fused_flow_reproduces_the_oracle(ode, mode, nseg, control):
    ph = W.make_phase(ode, mode, nseg, control_mode=control)
    check(ph.transcribe(), ph, (ode, mode))


@pytest.mark.parametrize("ode,mode,nseg", [("lowthrust", "LGL5", 44), ("lowthrust", "LGL3", 48), ("vanderpol", "LGL7", 40),
                                           ("lowthrust", "Trapezoidal", 48)])
def test_block_constant_control_through_the_fused_flow(ode, mode, nseg, monkeypatch):
    """ASSET_B200_FUSED_BLOCKED=1: the per-segment controls as satellite variables (fused.cpp: Finder::satellites)."""
    monkeypatch.setenv("ASSET_B200_FUSED_BLOCKED", "1")
    ph = W.make_phase(ode, mode, nseg, control_mode="BlockConstant")
    nlp = ph.transcribe()
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    check(nlp, ph, (ode, mode, "block-constant"))


@pytest.mark.parametrize("mode,name", [(4, "RHS"), (6, "OCC"), (1, "KKTNO"), (3, "AUG")])
def test_block_constant_control_other_modes_direct_outputs(mode, name, monkeypatch):
    """The residual-only and partial-matrix kernels of a block-constant phase: what the threads write directly against the
    oracle in that mode (their staged outputs follow the same lists as evalKKT's, checked above)."""
    monkeypatch.setenv("ASSET_B200_FUSED_BLOCKED", "1")
    ph = W.make_phase("lowthrust", "LGL5", 44, control_mode="BlockConstant")
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    R = info["t_hi"] - info["t_lo"]
    ref = oracle.Oracle.from_problem(nlp).eval(name, 0.7, X, LE, LI, sc)
    res = emulate_fused(nlp, X, LE, LI, sc, 0.7, [0, 1, 2, 3, R - 1], residual_mode=mode)
    res.pop("stage")
    wrote = 0
    for k, (v, m) in res.items():
        if m.any():
            wrote += int(m.sum())
            scale = max(float(np.abs(ref[k]).max()), 1e-300)
            assert float(np.abs(v[m] - ref[k][m]).max()) <= 1e-12 * scale, (name, k)
    assert wrote > 0


@pytest.mark.parametrize("mode", [1, 2, 3])
@pytest.mark.parametrize("ode,tm,nseg,control", [("lowthrust", "LGL5", 44, None), ("vanderpol", "LGL3", 48, None)])
def test_partial_matrix_modes_through_the_fused_flow(ode, tm, nseg, control, mode):
    """evalKKTNO / evalSOE / evalAUG with their fused kernels (ASSET_B200_FUSED_PARTIAL): zero-filled value array and staging
    slots, the mode's fused kernels, the mode's per-function programs at the mesh ends, the same gather lists."""
    ph = W.make_phase(ode, tm, nseg, control_mode=control)
    check(ph.transcribe(), ph, (ode, tm), mode)


@pytest.mark.parametrize("mode,control,nseg", [("LGL5", None, 44), ("LGL7", "HighestOrderSpline", 40), ("LGL3", None, 48),
                                               ("LGL5", "BlockConstant", 44), ("Trapezoidal", "BlockConstant", 48)])
def test_every_region_flag_with_parameters_through_the_fused_flow(mode, control, nseg, monkeypatch):
    """A phase with an ODE parameter, static parameters and functions over every region flag (helpers.regions_phase, without
    its accumulation, which is not fusable by design): rows of parameters and of the end times are far rows of every
    thread; with block-constant control the satellites join them."""
    from helpers import regions_phase
    if control == "BlockConstant":
        monkeypatch.setenv("ASSET_B200_FUSED_BLOCKED", "1")
    ph = regions_phase(mode, control, nseg, accumulate=False)
    nlp = ph.transcribe()
    assert nlp.fused_info()["ok"], nlp.fused_info()["why"]
    check(nlp, ph, ("regions", mode, control))
