import sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from asset_asrl_b200 import workloads as W
from helpers import problem_inputs
from emu_harness import emulate_fused
import oracle
mode, n = sys.argv[1], int(sys.argv[2])
ph = W.make_phase("lowthrust", mode, n); nlp = ph.transcribe(); X = ph.makeSolverInput(); LE, LI, sc = problem_inputs(nlp)
ref = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc)
t = time.time()
res = emulate_fused(nlp, X, LE, LI, sc, 0.7, [0, 1, 2, 3, 40])
print("emulated in %.1fs" % (time.time() - t))
for k, (v, m) in res.items():
    r = ref[k]
    scale = max(np.abs(r).max(), 1e-300)
    err = np.abs(v[m] - r[m]).max() / scale if m.any() else 0.0
    bad = np.nonzero(m)[0][np.abs(v[m] - r[m]) > 1e-12 * scale]
    print(k, "written", int(m.sum()), "max rel err", err, "bad idx", bad[:10])
