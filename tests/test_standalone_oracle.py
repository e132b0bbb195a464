"""Two independent builds of the bench workloads must agree: oracle/standalone.py (layout and index maps from the
independent restatement oracle/indexing.py, functions written down as s-expressions, no product code) against the
oracle mirrored from the product's transcription (Oracle.from_problem).  Sparsity pattern and KKTLocations bit-exact,
the solver input bit-exact, every output of evalKKT to rounding."""
import numpy as np
import pytest

import oracle
from oracle import standalone
from helpers import rel_err
from asset_asrl_b200 import workloads as W


@pytest.mark.parametrize("name,mode,control", [("lowthrust", "LGL5", None), ("lowthrust", "LGL7", "HighestOrderSpline"),
                                               ("lowthrust", "LGL3", None), ("lowthrust", "Trapezoidal", None),
                                               ("vanderpol", "LGL5", "NoSpline"), ("hypersens", "LGL7", None),
                                               ("brachistochrone", "LGL3", None), ("vanderpol", "Trapezoidal", None)])
@pytest.mark.parametrize("extras", [True, False])
def test_standalone_build_agrees_with_the_products_transcription(name, mode, control, extras):
    nseg = 23
    ph = W.make_phase(name, mode, nseg, control_mode=control, extras=extras)
    nlp = ph.transcribe()
    d = nlp.dims()
    mirrored = oracle.Oracle.from_problem(nlp)
    alone, X, dims = standalone.build(name, mode, nseg, control_mode=control, extras=extras)
    assert (dims["nvars"], dims["neq"], dims["niq"]) == (d["nvars"], d["neq"], d["niq"])
    assert np.array_equal(X, ph.makeSolverInput())
    pa, pm = alone.pattern(), mirrored.pattern()
    for a_, m_ in zip(pa, pm):
        assert np.array_equal(a_, m_)
    LE, LI = standalone.multipliers(dims)
    LE2, LI2 = W.multipliers(nlp)
    assert np.array_equal(LE, LE2) and np.array_equal(LI, LI2)
    sc = np.linspace(0.5, 1.5, alone.dims()["n_solver"])
    ra = alone.eval("KKT", 0.7, X, LE, LI, sc)
    rm = mirrored.eval("KKT", 0.7, X, LE, LI, sc)
    for k in ("KKT", "FXE", "FXI", "PGX", "AGX"):
        assert rel_err(ra[k], rm[k]) < 1e-14, k
    assert abs(ra["val"] - rm["val"]) <= 1e-14 * max(abs(rm["val"]), 1.0)
