"""Host pipeline without a GPU: expression -> scalar DAG -> symbolic adjoint -> device tape, executed by the numpy tape
interpreter (tests/tape_ref.py, same dual-number semantics as kernels.cu) and compared with the CPU oracle, function by
function and as assembled KKT outputs.  Also pins index maps / sparsity pattern against the oracle's own analysis."""
import json
import os

import numpy as np
import pytest

import oracle
from helpers import expression_zoo, interp_zoo, matrix_zoo, phase_zoo, table_phase_zoo, problem_inputs, rel_err, TOL
from ref_assembly import assemble
from prog_ref import run_program, dense_blocks
from asset_asrl_b200.Solvers import NonLinearProgram

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "collocation_functions.json")))
ZOO = phase_zoo()
ZOO.update(table_phase_zoo())


@pytest.mark.parametrize("name,fun,gen", expression_zoo() + matrix_zoo() + interp_zoo(), ids=[z[0] for z in expression_zoo() + matrix_zoo() + interp_zoo()])
def test_tape_matches_oracle(name, fun, gen):
    rng = np.random.default_rng(5)
    ir, orr = fun.IRows(), fun.ORows()
    nlp = NonLinearProgram(ir, orr, 0)
    t = nlp.add_tape(fun, True)
    nlp.add_function(1, t, False, np.arange(ir, dtype=np.int32).reshape(1, ir), np.arange(orr, dtype=np.int32).reshape(1, orr))
    fi = nlp.function_info(0)
    for mode in (4, 3, 1, 0):
        st, code, consts = nlp.program(0, mode)
        tables = nlp.program_tables(0, mode)
        for _ in range(3):
            x, lam = gen(rng), rng.uniform(-1, 1, orr)
            f, J, g, H = dense_blocks(fi, *run_program(st, code, consts, x, lam, tables=tables))
            fx, jx, gx, hx = oracle.computeall("expr " + fun.describe(), x, lam)
            assert rel_err(f, fx) < TOL
            if mode >= 3:
                assert rel_err(J, jx) < TOL
            if mode in (1, 3, 4):
                assert rel_err(g, gx) < TOL
            if mode == 4:
                assert rel_err(H, hx) < TOL


@pytest.mark.parametrize("name", sorted(ZOO), ids=sorted(ZOO))
def test_assembled_kkt_and_pattern(name):
    ph = ZOO[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    orc = oracle.Oracle.from_problem(nlp)
    # sparsity pattern and KKTLocations: bit-exact
    ip, ix, loc = orc.pattern()
    ip2, ix2 = nlp.pattern()
    assert np.array_equal(ip, ip2) and np.array_equal(ix, ix2) and np.array_equal(loc, nlp.kkt_locations())
    ref = orc.eval("KKT", 0.7, X, LE, LI, sc)
    out = assemble(nlp, X, LE, LI, 0.7, solver_coeffs=sc)
    for k in ("KKT", "FXE", "FXI", "PGX", "AGX"):
        assert rel_err(out[k], ref[k]) < TOL, k
    assert abs(out["val"] - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0)


def test_tape_matches_golden_defects():
    """The lowered defect tapes against the 50-digit fixtures (independent of the oracle)."""
    from test_oracle import _golden_desc  # noqa: F401  (desc strings are also valid product descriptions)
    from helpers import vdp_param_ode
    from asset_asrl_b200 import workloads as W
    odes = {"lowthrust": W.lowthrust_ode, "brachistochrone": W.brachistochrone_ode, "vanderpol_mu": vdp_param_ode}
    for c in GOLDEN:
        if c["kind"] != "defect":
            continue
        ode = odes[c["ode"]]()
        S = c["CS"]
        n = c["XV"] + 1 + c["UV"]
        x = np.array(c["x"])
        T = np.zeros((S, n + c["PV"]))
        T[:, :n] = x[:S * n].reshape(S, n)
        if c["PV"]:
            T[:, n:] = x[S * n:]
        ph = ode.phase(c["mode"], T, 1)
        nlp = ph.transcribe()
        fi = nlp.function_info(0)
        st, code, consts = nlp.program(0, 4)
        f, J, g, H = dense_blocks(fi, *run_program(st, code, consts, x, np.array(c["lam"])))
        assert rel_err(f, c["fx"]) < TOL and rel_err(J, c["jx"]) < TOL and rel_err(H, c["hx"]) < TOL, c["name"]


def test_linearity_rules():
    """Type-level linearity decides whether Hessian triplets exist (note L): one assert per row of
    the rule table.  A function built by a Python operator is a GenericFunction (static IsLinearFunction false,
    ComputableBase.h:90); the solver unwraps ONE level (SolverInterfaceSpecs.h:386-388), so an operator applied to
    Arguments / Segments (concrete linear types) is linear, the same operator applied to an operator result is not."""
    from helpers import vf, oc
    a = vf.Arguments(3)
    lin_of_args = {
        "Arguments (Segment.h:102)": a,
        "Segment": a.head(2),
        "Elements (Elements.h:21)": a[[0, 2]],
        "coeff": a[1],
        "FunctionDifference<Segment,Segment> (Summation.h:132)": a[1] - a[0],
        "TwoFunctionSum": a[1] + a[0],
        "MultiFunctionSum (Summation.h:349)": vf.sum(a[0], a[1], a[2]),
        "Scaled<Arguments> (Scaled.h:81)": a * 2.0,
        "Scaled by division": a / 4.0,
        "negation": -a,
        "RowScaled (Scaled.h:320)": a * np.array([1.0, 2.0, 3.0]),
        "FunctionPlusVector (FunctionVectorSums.h:30)": a + np.array([1.0, 2.0, 3.0]),
        "VectorMinusFunction": np.array([1.0, 2.0, 3.0]) - a,
        "function minus vector": a - np.array([1.0, 2.0, 3.0]),
        "StackedOutputs of segments (Stacked.h:85)": vf.stack([a[0], a[1]]),
        "CwiseSum (CwiseSum.h:28)": a.sum(),
        "MatrixScaled (Scaled.h:483)": vf.matmul(np.arange(6.0).reshape(2, 3), a),
        "IOScaled<Arguments> (IOScaled.h:14)": vf.ioscaled(a, [1.0, 2.0, 3.0], [2.0, 2.0, 2.0]),
        "Constant (Constant.h:13)": vf.ConstantVector(3, [1.0, 2.0]),
    }
    for name, f in lin_of_args.items():
        assert f.is_linear(), name
    g = a[1] - a[0]           # a GenericFunction holding a linear type
    h = a.head(2) * 3.0       # likewise
    hidden = {
        "Scaled<GenericFunction>": g * 2.0,
        "FunctionPlusVector<GenericFunction>": h + np.array([1.0, 2.0]),
        "VectorMinusFunction<GenericFunction>": np.array([1.0, 2.0]) - h,
        "TwoFunctionSum<Generic,Generic>": g + g,
        "FunctionDifference<Generic,Segment>": g - a[2],
        "StackedOutputs<Generic,Segment> (all parts must be linear types)": vf.stack([g, a[2]]),
        "Segment of a GenericFunction": h[0],
        "CwiseSum<GenericFunction>": h.sum(),
        "RowScaled<GenericFunction>": h * np.array([1.0, 2.0]),
        "MatrixScaled<GenericFunction>": vf.matmul(np.eye(2), h),
        "IOScaled<GenericFunction>": vf.ioscaled(h, [1.0, 1.0, 1.0], [2.0, 2.0]),
        "NestedFunction<Generic,Generic> (outer && inner, NestedFunction.h:38)": h.eval(a * 2.0),
    }
    for name, f in hidden.items():
        assert not f.is_linear(), name
    nonlinear = {
        "sin": vf.sin(a), "cos": vf.cos(a), "exp": vf.exp(a[0]), "squared": a[0].squared(), "sqrt": vf.sqrt(a[0]),
        "inverse": a[0].inverse(), "pow": a[0] ** 3, "abs": abs(a[0]), "sign": a[0].sign(),
        "norm": a.norm(), "squared_norm": a.squared_norm(), "inverse_norm": a.inverse_norm(), "normalized": a.normalized(),
        "dot": a.dot(a), "cross": a.cross(a), "cwiseProduct": a.cwiseProduct(a), "cwiseQuotient": a.cwiseQuotient(a),
        "scalar product of functions": a[0] * a[1], "function quotient": a[0] / a[1], "vector times scalar function": a * a[0],
        "arctan2": vf.arctan2(a[0], a[1]), "ifelse": vf.ifelse(a[0] < 0.0, a[1], a[2]),
    }
    for name, f in nonlinear.items():
        assert not f.is_linear(), name
    # helpers the reference builds in C++ keep their concrete type: addLUVarBound is
    # StackedOutputs<Scaled<FunctionPlusVector<Scaled<Arguments<1>>>>, Scaled<FunctionPlusVector<Arguments<1>>>>
    # (ODEPhaseBase.cpp:110-130), linear; the norm bound composes it with a norm (:226-260), non-linear
    assert oc._lu_bound(-1.0, 1.0, 2.0, 2.0).is_linear()
    ph = __import__("asset_asrl_b200").workloads.make_phase("lowthrust", "LGL3", 4)
    nlp = ph.transcribe()
    lin = {nlp.function_info(f)["linear"] for f in range(nlp.dims()["nfuncs"])}
    assert lin == {0, 1}

def test_link_index_maps_follow_the_reference_layout():
    """make_link_Vindex_Cindex (OptimalControlProblem.cpp:659-764) on a hand-computed case: phase variables are
    [(x,t,u) per node ; p], phases are laid out back to back, link parameters follow the last phase, link constraint
    rows follow every phase's rows in insertion order, application-major."""
    ocp = ZOO["ocp_forward_links_mixed_transcriptions"]()
    nlp = ocp.transcribe()
    nv = [5 * 4 + 1, 7 * 4 + 1, 6 * 4 + 1]  # LGL3/4 segments, LGL5/3, Trapezoidal/5; XtU = 4, PV = 1
    off = np.concatenate([[0], np.cumsum(nv)])
    assert [o[0] for o in ocp.offsets] == list(off[:3])
    d = nlp.dims()
    assert d["nvars"] == off[3]
    back = lambda p, v: off[p] + (nv[p] - 1) - 4 + v  # noqa: E731  last node's variable v
    front = lambda p, v: off[p] + v  # noqa: E731
    par = lambda p: off[p] + nv[p] - 1  # noqa: E731
    links = []
    for f in range(d["nfuncs"]):
        fi = nlp.function_info(f)
        V, Cx = nlp.function_indices(f)
        links.append((fi, V, Cx))
    eq_links = [(fi, V, Cx) for (fi, V, Cx) in links if fi["role"] == 1][-4:]
    # addForwardLinkEqualCon(0, -1, [0,1,2]) -> two functions; direct ODEParams link; flag-interface link with two columns
    assert np.array_equal(eq_links[0][1], [[back(0, 0), back(0, 1), back(0, 2), front(1, 0), front(1, 1), front(1, 2)]])
    assert np.array_equal(eq_links[1][1], [[back(1, 0), back(1, 1), back(1, 2), front(2, 0), front(2, 1), front(2, 2)]])
    assert np.array_equal(eq_links[2][1], [[par(0), par(1)]])
    assert np.array_equal(eq_links[3][1], [[back(0, 3), front(1, 3)], [back(1, 3), front(2, 3)]])
    rows = np.concatenate([c.ravel() for (_, _, c) in eq_links[:4]])
    assert np.array_equal(rows, np.arange(d["neq"] - rows.size, d["neq"])) and rows.size == 3 + 3 + 1 + 2
    iq = [(fi, V, Cx) for (fi, V, Cx) in links if fi["role"] == 2][-1]
    assert np.array_equal(iq[1], [[front(0, 0), front(0, 1), back(2, 0), back(2, 1)]]) and iq[2][0, 0] == d["niq"] - 1
    ob = [(fi, V, Cx) for (fi, V, Cx) in links if fi["role"] == 0][-1]
    assert np.array_equal(ob[1], [[front(0, 2), back(2, 2)]])
    # link parameters: after the last phase
    ocp2 = ZOO["ocp_multispacecraft_LGL5_block_constant"]()
    nlp2 = ocp2.transcribe()
    d2 = nlp2.dims()
    assert np.array_equal(ocp2.LinkParamLocs, d2["nvars"] - 7 + np.arange(7))
    assert ocp2.makeSolverInput().shape == (d2["nvars"],)
    fi, V, Cx = [(nlp2.function_info(f),) + tuple(nlp2.function_indices(f)) for f in range(d2["nfuncs"])
                 if nlp2.function_info(f)["role"] == 1][-1]
    assert np.array_equal(V, [ocp2.LinkParamLocs[:6]]) and Cx[0, 0] == d2["neq"] - 1
