"""Dry run of bench.py's GPU arm on a CPU-only machine, N ranks over gloo: torch.cuda and the device plan are mocked (values
from the oracle, one dummy all-reduce per sharded evaluation), so bench.py's own control flow and collectives run.  The numbers it prints mean nothing.  Usage (test infrastructure):
  python tools/bench_dry_run.py --steps 3 --warmup 1 --segments 25000
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
      tools/bench_dry_run.py --gpus 2 --steps 3 --warmup 1 --segments 25000 --no-cpu --no-parity   [--shard one-phase|phases]
The mocked clock sampler delivers its lines to rank r only after r + 1 looks, which is what exposed the per-rank exit of
the warm loop (ranks leaving after different numbers of collectives)."""
import os, sys, time, types
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import torch.distributed as dist

RANK = int(os.environ.get("RANK", 0))
# ---- torch.cuda / NCCL stand-ins
torch.cuda.is_available = lambda: True
torch.cuda.set_device = lambda d: None
torch.cuda.synchronize = lambda *a, **k: None
class _Ev:
    def __init__(self, **k): self.t = 0.0
    def record(self, s=None): self.t = time.perf_counter()
    def elapsed_time(self, o): return (o.t - self.t) * 1e3 + 1e-3
torch.cuda.Event = _Ev
torch.cuda.ExternalStream = lambda *a, **k: None
_tensor, _zeros = torch.tensor, torch.zeros
def _strip(f):
    def g(*a, **k):
        if str(k.get("device", "")).startswith("cuda"): k.pop("device")
        return f(*a, **k)
    return g
torch.tensor, torch.zeros = _strip(_tensor), _strip(_zeros)
torch.Tensor.cuda = lambda self, *a, **k: self
torch.Tensor.pin_memory = lambda self, *a, **k: self
_init = dist.init_process_group
dist.init_process_group = lambda backend, **k: _init("gloo")

import oracle
from asset_asrl_b200 import Solvers as S, _lib
NAMES = ["KKT", "KKTNO", "SOE", "AUG", "RHS", "OGC", "OCC", "OBJ"]
P = S.NonLinearProgram
def to_device(self, device=0, rank=0, world=1, group=None):
    self.plan, self.device, self.rank, self.world, self.group = 1, device, rank, world, group
    self._orc, self._cache, self._sc = oracle.Oracle.from_problem(self), {}, None
    return self
def _res(self, mode, scale):
    key = (mode, scale)
    if key not in self._cache:
        self._cache[key] = self._orc.eval(NAMES[mode], scale, self._X, self._LE, self._LI, self._sc)
    return self._cache[key]
def _collective(self):
    if self.world > 1:
        dist.all_reduce(torch.zeros(7, dtype=torch.float64))
def eval_device(self, mode, scale):
    self._last = _res(self, mode, float(scale)); _collective(self)
def set_inputs(self, X, LE, LI): self._X, self._LE, self._LI, self._cache = X, LE, LI, {}
def device_tensor(self, name):
    if name in ("X", "LE", "LI"): return torch.zeros(4, dtype=torch.float64)
    v = self._last[name]; return torch.from_numpy(np.atleast_1d(np.asarray(v, dtype=np.float64)).copy())
def output_ranges(self):
    d = self.dims(); return {"val": [(0, 1)], "PGX": [(0, d["nvars"])], "AGX": [(0, d["nvars"])], "FXE": [(0, d["neq"])], "FXI": [(0, d["niq"])], "KKT": [(0, d["nnz"])]}
def eval_host(self, mode, scale, X, LE, LI, val, PGX, AGX, FXE, FXI, KKT, **k):
    set_inputs(self, X, LE, LI); r = _res(self, mode, float(scale)); _collective(self)
    for arr, nm in ((val, "val"), (PGX, "PGX"), (AGX, "AGX"), (FXE, "FXE"), (FXI, "FXI"), (KKT, "KKT")):
        if arr is not None:
            v = np.atleast_1d(np.asarray(r[nm], dtype=np.float64)); n = min(arr.size, v.size); arr[:n] = v[:n]
def input_ranges(self):
    saved, self.plan = self.plan, None
    try: return P._input_ranges_real(self)
    finally: self.plan = saved
P._input_ranges_real = P.input_ranges
for nm, fn in dict(to_device=to_device, eval_device=eval_device, set_inputs=set_inputs, device_tensor=device_tensor, output_ranges=output_ranges,
                   eval_host=eval_host, eval_host_sharded=eval_host, input_ranges=input_ranges).items():
    setattr(P, nm, fn)
P.set_solver_coeffs = lambda self, c: setattr(self, "_sc", np.asarray(c, dtype=float).copy())
P.sync = lambda self: None
P.stream = lambda self: 0
P.launches = lambda self: 21
P.specialised_launches = lambda self: 20
P.profile = lambda self, *a: None
P.profile_times = lambda self: [("fused", 0.4), ("gather:KKT", 0.01)]
P.fused_segments = lambda self: 1000
P.__del__ = lambda self: None
_lib.lib.ab_plan_set_inputs_sliced = lambda *a: 0
_lib.lib.ab_plan_get_outputs_packed = lambda *a: 0

import bench
bench._lib_cut_count = lambda nlp: 7
class _Clk:
    def __init__(self, dev): self.lines, self.n = [], 0
    def start(self): self.t0 = time.time()
    @property
    def _(self): return None
    def stop(self, w0, w1): return {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "reasons": [], "samples": 3}
class _Lines(list):
    # rank-dependent delivery: rank r sees three lines only after r + 1 looks, so the ranks would leave the warm loop at
    # different times if the exit were decided locally
    def __init__(self): super().__init__(); self.looks = 0
    def __len__(self):
        self.looks += 1
        return 3 if self.looks > RANK + 1 else 0
_Clk.lines = property(lambda self: self._l, lambda self, v: setattr(self, "_l", _Lines()))
bench.ClockSampler = _Clk
sys.argv = ["bench.py"] + sys.argv[1:]
a = bench.parse_args()
bench.run_b200(a)
print("DRY RUN COMPLETE rank", RANK, flush=True)
