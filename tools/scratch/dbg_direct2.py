import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from asset_asrl_b200 import workloads as W
from helpers import problem_inputs
import oracle
ph = W.make_phase("lowthrust", "LGL7", 90, control_mode="HighestOrderSpline"); nlp = ph.transcribe(); X = ph.makeSolverInput(); LE, LI, sc = problem_inputs(nlp)
nlp.to_device(0); nlp.set_solver_coeffs(sc)
out = nlp.evalKKT(0.7, X, LE, LI)
ref = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc)
for k in ("PGX", "AGX", "FXE", "FXI", "KKT"):
    a, b = ref[k], out[k]
    bad = np.nonzero(np.abs(a - b) > 1e-12 * max(1.0, np.abs(a).max()))[0]
    print(k, a.size, "mismatches", bad.size, bad[:12])
print("ref PGX", ref["PGX"][120:160]); print("out PGX", out["PGX"][120:160]); print("out AGX", out["AGX"][120:160])
