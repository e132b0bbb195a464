"""Oracle windows (oracle/windows.py): the sub-problem built from the applications touching a variable window reproduces
the full oracle inside the window -- checked here on the CPU with the full oracle standing in for the device output."""
import numpy as np
import pytest

import oracle
from oracle.windows import check_windows
from helpers import phase_zoo, problem_inputs
from asset_asrl_b200 import workloads as W

ZOO = phase_zoo()


@pytest.mark.parametrize("name", ["lowthrust_300", "regions_LGL5", "regions_LGL5_block_constant", "lowthrust_LGL7"])
def test_windows_reproduce_the_full_oracle(name):
    ph = W.make_phase("lowthrust", "LGL5", 300) if name == "lowthrust_300" else ZOO[name]()
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    full = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc)
    nv = nlp.dims()["nvars"]
    windows = [(0, min(40, nv)), (nv // 2, min(nv // 2 + 37, nv)), (max(nv - 45, 0), nv)]
    worst = check_windows(nlp, full, windows, 0.7, X, LE, LI, sc)
    assert worst["rows"] == sum(b - a for a, b in windows)
    for k in ("KKT", "AGX", "PGX", "FXE", "FXI"):
        if k in worst:
            assert worst[k] < 1e-13, (k, worst)


def test_a_wrong_value_inside_a_window_is_seen():
    ph = W.make_phase("lowthrust", "LGL5", 120)
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    full = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc)
    ip, ix = nlp.pattern()
    bad = {k: (np.array(v, copy=True) if isinstance(v, np.ndarray) else v) for k, v in full.items()}
    seg = slice(int(ip[1200]), int(ip[1210]))
    bad["KKT"][seg.start + int(np.argmax(np.abs(bad["KKT"][seg])))] *= 1.0 + 1e-9
    assert check_windows(nlp, bad, [(1200, 1210)], 0.7, X, LE, LI, sc)["KKT"] > 1e-12
    assert check_windows(nlp, bad, [(100, 140)], 0.7, X, LE, LI, sc)["KKT"] < 1e-13
