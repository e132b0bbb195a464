"""Workload for compute-sanitizer (tools/sanitize.sh): every kind of launch the evaluation makes, on meshes small enough
for the tool's slowdown -- fused evalKKT (graph replay included), the seven other modes on the per-function kernels,
a sharded and a windowed plan, block-constant control (interpreter-free per-function path) -- each checked against the
oracle so that a run that 'passes' the tool also computed the right thing."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from asset_asrl_b200 import workloads as W  # noqa: E402
from asset_asrl_b200.Solvers import EVAL_KKT  # noqa: E402
from helpers import problem_inputs, rel_err  # noqa: E402

MODES = {"KKT": ("PGX", "AGX", "FXE", "FXI", "KKT"), "KKTNO": ("AGX", "FXE", "FXI", "KKT"), "SOE": ("FXE", "FXI", "KKT"),
         "AUG": ("PGX", "AGX", "FXE", "FXI", "KKT"), "RHS": ("PGX", "AGX", "FXE", "FXI")}
nseg = int(sys.argv[1]) if len(sys.argv) > 1 else 600
worst = 0.0
for mode, cm in (("LGL5", None), ("LGL7", None), ("LGL5", "BlockConstant")):
    ph = W.make_phase("lowthrust", mode, nseg, control_mode=cm)
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    orc = oracle.Oracle.from_problem(nlp)
    nlp.to_device(0)
    nlp.set_solver_coeffs(sc)
    calls = {"KKT": nlp.evalKKT, "KKTNO": nlp.evalKKTNO, "SOE": nlp.evalSOE, "AUG": nlp.evalAUG, "RHS": nlp.evalRHS}
    for rep in range(3):  # the third call of a mode replays its graph
        for name, fn in calls.items():
            out = fn(0.7, X, LE, LI)
            ref = orc.eval(name, 0.7, X, LE, LI, sc)
            for k in MODES[name]:
                worst = max(worst, rel_err(out[k], ref[k]))
    print(mode, cm, "fused segments", nlp.fused_segments(), "worst", worst, flush=True)
# windowed plans, world 3 on one device
ph = W.make_phase("lowthrust", "LGL5", nseg)
whole = ph.transcribe()
X = ph.makeSolverInput()
LE, LI, sc = problem_inputs(whole)
ref = oracle.Oracle.from_problem(whole).eval("KKT", 0.7, X, LE, LI, sc)
agx = np.zeros_like(ref["AGX"])
for r in range(3):
    nlp = ph.transcribe_window(r, 3)
    nlp.to_device(0, rank=r, world=3)
    nlp.set_solver_coeffs(sc)
    nlp.set_inputs(X, LE, LI)
    nlp.world = 1
    for _ in range(3):
        nlp.eval_device(EVAL_KKT, 0.7)
    nlp.sync()
    o = nlp.get_outputs()
    fe = o["FXE"]
    own = fe != 0
    worst = max(worst, float(np.abs(fe[own] - ref["FXE"][own]).max()) / float(np.abs(ref["FXE"]).max()))
print("windowed worst", worst, flush=True)
assert worst < 1e-12, worst
print("sanitize workload ok")
