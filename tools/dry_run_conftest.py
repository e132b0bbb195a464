"""Dry run of GPU test files on a machine without a GPU: the device plan is replaced by the CPU oracle, so only the tests' own
Python (imports, fixtures, keys, mode tables, environment switches) is exercised -- it says nothing about the kernels.
Usage: mkdir /tmp/dry && cp tools/dry_run_conftest.py /tmp/dry/conftest.py && cp tests/test_zz2_gpu_tables.py /tmp/dry/ &&
       (cd /tmp/dry && python -m pytest -x -q -p no:cacheprovider .)      (test infrastructure; never imported by the product)"""
import numpy as np, pytest, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import oracle
from asset_asrl_b200 import Solvers as S

NAMES = ["KKT", "KKTNO", "SOE", "AUG", "RHS", "OGC", "OCC", "OBJ"]
def to_device(self, device=0, rank=0, world=1, group=None):
    self.plan = 1; self.device = device; self.rank, self.world, self.group = rank, world, group; self._sc = None
    self._orc = oracle.Oracle.from_problem(self); return self
def set_solver_coeffs(self, c): self._sc = np.asarray(c, dtype=float).copy()
def _eval(self, mode, ObjScale, X, LE, LI, want):
    d = self.dims()
    LE = LE if LE is not None else np.zeros(d["neq"]); LI = LI if LI is not None else np.zeros(d["niq"])
    r = self._orc.eval(NAMES[mode], ObjScale, X, LE, LI, self._sc)
    self._last = mode
    out = {}
    if "val" in want: out["val"] = float(r["val"])
    for k in ("PGX", "AGX", "KKT"):
        if k in want: out[k] = r[k]
    if "FX" in want: out["FXE"], out["FXI"] = r["FXE"], r["FXI"]
    return out
S.NonLinearProgram.to_device = to_device
S.NonLinearProgram.set_solver_coeffs = set_solver_coeffs
S.NonLinearProgram._eval = _eval
S.NonLinearProgram.fused_segments = lambda self: 1 if (self.fused_info()["ok"] and (self._last in (0, 4, 5, 6, 7) or __import__("os").environ.get("ASSET_B200_FUSED_PARTIAL"))) else 0
S.NonLinearProgram.specialised_launches = lambda self: 0 if __import__("os").environ.get("ASSET_B200_JIT") == "0" else 1
S.NonLinearProgram.__del__ = lambda self: None
