"""Rank-local analysis (Problem::window, ODEPhase.transcribe_window): every rank analyses only its own mesh segments
plus a halo, and the assembled result must be the whole problem's -- checked against the oracle evaluated on the whole
problem, row by row through both CSR patterns."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import oracle  # noqa: E402
from helpers import problem_inputs, rel_err, TOL  # noqa: E402

from asset_asrl_b200 import workloads as W  # noqa: E402
from asset_asrl_b200.Solvers import EVAL_KKT  # noqa: E402

pytestmark = pytest.mark.gpu


def run_windowed(ph, world, X, LE, LI, sc, obj_scale=0.7):
    import torch
    plans = []
    for r in range(world):
        nlp = ph.transcribe_window(r, world)
        nlp.to_device(0, rank=r, world=world)
        nlp.set_solver_coeffs(sc)
        nlp.set_inputs(X, LE, LI)
        plans.append(nlp)
    for p in plans:
        p.world = 1  # no NCCL in this test: the cut set is summed by hand below
        p.eval_device(EVAL_KKT, obj_scale)
        p.world = world
        p.cut_pack()
        p.sync()
    total = sum(p.cut_tensor().clone() for p in plans)
    for p in plans:
        p.cut_tensor().copy_(total)
        torch.cuda.synchronize()
        p.cut_unpack()
        p.sync()
    return plans, [p.get_outputs() for p in plans]


@pytest.mark.parametrize("mode,nseg,world", [("LGL5", 2000, 4), ("LGL3", 900, 3), ("LGL7", 700, 2), ("LGL5", 1300, 8)])
def test_windowed_plans_assemble_the_whole_problem(mode, nseg, world):
    ph = W.make_phase("lowthrust", mode, nseg)
    whole = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(whole)
    ref = oracle.Oracle.from_problem(whole).eval("KKT", 0.7, X, LE, LI, sc, threads=oracle.hardware_threads())
    gip, gix = whole.pattern()
    d = whole.dims()
    plans, outs = run_windowed(ph, world, X, LE, LI, sc)
    # analysis cost follows the share, not the phase
    for p in plans:
        assert p.dims()["n_user"] < 1.5 * d["n_user"] / world + 40000
        assert p.fused_segments() > 0
    for k in ("PGX", "AGX", "FXE", "FXI"):
        stack = np.stack([o[k] for o in outs])
        assert np.all((stack != 0).sum(axis=0) <= 1), (k, "an output is held by more than one rank")
        assert rel_err(stack.sum(axis=0), ref[k]) < TOL, k
    assert abs(sum(o["val"] for o in outs) - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0)
    # KKT values: rows of the whole pattern rebuilt from the ranks' own patterns
    got = np.zeros(d["nnz"])
    holders = np.zeros(d["nnz"], dtype=np.int32)
    for p, o in zip(plans, outs):
        lip, lix = p.pattern()
        far = set(int(v) for v in p.window_far_vars())
        for r in range(d["kktdim"]):
            a0, a1, b0, b1 = int(gip[r]), int(gip[r + 1]), int(lip[r]), int(lip[r + 1])
            v = o["KKT"][b0:b1]
            if not np.any(v != 0):
                continue
            if a1 - a0 == b1 - b0:
                assert np.array_equal(gix[a0:a1], lix[b0:b1]), ("row pattern differs", r)
                pos = np.arange(a0, a1)
            else:
                # rows of the variables every application of a function reads (the phase's first and last time): a rank
                # holds the columns of its own applications, a subset of the whole row
                assert r in far, ("row pattern differs", r)
                pos = a0 + np.searchsorted(gix[a0:a1], lix[b0:b1])
                assert np.array_equal(gix[pos], lix[b0:b1]), ("columns missing from the whole row", r)
            got[pos] += v
            holders[pos] += (v != 0)
    assert holders.max() <= 1, "a KKT entry is held by more than one rank"
    assert rel_err(got, ref["KKT"]) < TOL
