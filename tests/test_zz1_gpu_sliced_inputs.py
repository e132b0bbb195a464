"""Sliced upload of the iterate for windowed plans (ab_plan_set_inputs_sliced): a rank receives
only the ranges of X / LE / LI its functions read.  With the device copy of the iterate poisoned by NaN first, a sliced
evaluation must reproduce a full-upload evaluation bit for bit on every rank's plan.

A file of its own, collected late: written when the round's GPU time was spent (the ranges themselves are tested on the
CPU in tests/test_window_host.py)."""
import numpy as np
import pytest

from helpers import problem_inputs
from asset_asrl_b200 import workloads as W
from asset_asrl_b200.Solvers import EVAL_KKT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,nseg,world", [("LGL5", 4000, 8), ("LGL3", 900, 3)])
def test_sliced_upload_reproduces_the_full_upload(mode, nseg, world):
    import torch
    ph = W.make_phase("lowthrust", mode, nseg)
    whole = ph.transcribe()
    d = whole.dims()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(whole)
    for r in range(world):
        nlp = ph.transcribe_window(r, world)
        nlp.to_device(0, rank=r, world=world)
        nlp.set_solver_coeffs(sc)
        nlp.world = 1  # one plan at a time, no collective: the cut set plays no part in what a rank reads
        nlp.set_inputs(X, LE, LI)
        nlp.eval_device(EVAL_KKT, 0.7)
        want = nlp.get_outputs()
        rg = nlp.input_ranges()
        assert sum(hi - lo for lo, hi in rg["X"]) < 1.3 * d["nvars"] / world + 40 * d["nvars"] / nseg
        for nm in ("X", "LE", "LI"):
            t = nlp.device_tensor(nm)
            if t.numel():
                t.fill_(float("nan"))
        torch.cuda.synchronize()
        from asset_asrl_b200 import _lib
        _lib.check0(_lib.lib.ab_plan_set_inputs_sliced(nlp.plan, _lib.dptr(X), _lib.dptr(LE), _lib.dptr(LI)))
        nlp.eval_device(EVAL_KKT, 0.7)
        got = nlp.get_outputs()
        for k in ("PGX", "AGX", "FXE", "FXI", "KKT"):
            assert np.array_equal(got[k], want[k]), (r, k)
        assert got["val"] == want["val"]
