import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from asset_asrl_b200 import workloads as W
from helpers import problem_inputs
name = sys.argv[1] if len(sys.argv) > 1 else "hos"
mk = {"hos": lambda: W.make_phase("lowthrust", "LGL7", 90, control_mode="HighestOrderSpline"),
      "lgl7": lambda: W.make_phase("lowthrust", "LGL7", 100)}[name]
res = {}
for direct in ("0", "1"):
    os.environ["ASSET_B200_FUSED_DIRECT"] = direct
    ph = mk(); nlp = ph.transcribe(); X = ph.makeSolverInput(); LE, LI, sc = problem_inputs(nlp)
    info = nlp.fused_info(); print(direct, info)
    nlp.to_device(0); nlp.set_solver_coeffs(sc)
    res[direct] = nlp.evalKKT(0.7, X, LE, LI)
    print("fused segs", nlp.fused_segments())
for k in ("PGX", "AGX", "FXE", "FXI", "KKT"):
    a, b = res["0"][k], res["1"][k]
    bad = np.nonzero(np.abs(a - b) > 1e-12 * max(1.0, np.abs(a).max()))[0]
    print(k, a.size, "mismatches", bad.size, bad[:40], (a[bad[:8]], b[bad[:8]]) if bad.size else "")
