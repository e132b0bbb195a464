#!/usr/bin/env python
"""bench.py -- KKT evaluation throughput of the collocation NLP (constraints + Jacobian + Hessian + scatter).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU evaluation (oracle port) on the host cores

A step is one NonLinearProgram::evalKKT over every mesh segment of the workload phase.
``value`` = segments/s with X, LE, LI resident in HBM when the timed region starts; ``e2e`` = the same through the
C ABI's host-buffer call (H2D of X/LE/LI and D2H of every output inside the timed region).  The reference's own
binary cannot be built in this image, so the CPU numbers come from oracle/ (kind "port").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "kkt_eval_segments_per_sec"
UNIT = "segments/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ode", default="lowthrust", choices=["lowthrust", "vanderpol", "hypersens", "brachistochrone", "minclimb"])
    ap.add_argument("--mode", default="LGL5", choices=["LGL3", "LGL5", "LGL7", "Trapezoidal"])
    ap.add_argument("--segments", type=int, default=int(os.environ.get("AB_BENCH_SEGMENTS", 200000)),
                    help="mesh segments per GPU (weak scaling: each rank owns a contiguous range of the phase)")
    ap.add_argument("--cpu-segments", type=int, default=20000, help="segments of the bounded CPU sample")
    ap.add_argument("--no-extras", action="store_true", help="defect constraints only (no bounds / integral objective)")
    ap.add_argument("--control-mode", default=None, choices=["HighestOrderSpline", "FirstOrderSpline", "NoSpline", "BlockConstant"],
                    help="control parameterisation of the phase (default: the reference's, FirstOrderSpline)")
    ap.add_argument("--eval", default="KKT", choices=["KKT", "KKTNO", "SOE", "AUG", "RHS", "OGC", "OCC", "OBJ"],
                    help="which NonLinearProgram::eval* entry point a step is (default evalKKT, the headline metric; the parity "
                         "leg, the roofline bytes and e2e are evalKKT's and are skipped for the others)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--fused-partial", action="store_true",
                    help="--eval KKTNO / SOE / AUG on the fused threads (ASSET_B200_FUSED_PARTIAL=1; opt-in until run on a device)")
    ap.add_argument("--fused-blocked", action="store_true",
                    help="--control-mode BlockConstant on the fused threads (ASSET_B200_FUSED_BLOCKED=1; opt-in until run on a device)")
    ap.add_argument("--full-upload", action="store_true",
                    help="windows sharding: upload the whole iterate on every rank each step instead of the rank's slice")
    ap.add_argument("--no-parity", action="store_true", help="skip the (untimed) oracle-window comparison before the timed region")
    ap.add_argument("--shard", default="windows", choices=["windows", "one-phase", "phases"],
                    help="N > 1: one phase of N x segments split across the ranks -- 'windows' (default): every rank analyses "
                         "only its own segments plus a halo (rank-local analysis, no cap on the total); 'one-phase': every "
                         "rank analyses the whole problem and takes its share (capped by --max-total-segments); both sum "
                         "the cut set over NCCL -- or 'phases': one independent phase per rank (no collective)")
    ap.add_argument("--max-total-segments", type=int, default=800000,
                    help="one-phase sharding: every rank analyses the whole problem on the host (about 45 KB and 0.1 ms "
                         "per segment), so the total is capped and segments per GPU shrink beyond the cap")
    return ap.parse_args()


def workload_name(a):
    return (f"{a.ode}_{a.mode.lower()}" + ("_defect_only" if a.no_extras else "_example_functions")
            + (f"_{a.control_mode.lower()}" if getattr(a, "control_mode", None) else ""))


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        # the timed region is a few milliseconds: take the samples inside it, else those of the loaded half second around it
        rows = ([ln.split(",") for (t, ln) in self.lines if t0 - 0.02 <= t <= t1 + 0.02]
                or [ln.split(",") for (t, ln) in self.lines if t0 - 0.5 <= t <= t1 + 0.1] or [ln.split(",") for (_, ln) in self.lines])
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, nm in enumerate(names):
                    if r[3 + k].strip().lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_eval_rate(a, steps, warmup, full_size=False):
    """Oracle (CPU port of the reference evaluation) on a bounded sample of the same workload, all host threads: once
    with the IEEE build used for parity and once built with the reference's own flags (-O2 -march=native -ffast-math,
    CMakeLists.txt:55,:86,:160); the faster of the two is the reported value."""
    import oracle
    from oracle import standalone
    nseg = a.segments if full_size else min(a.cpu_segments, a.segments)
    # The problem is built without the product where the standalone builder covers the workload (oracle/standalone.py:
    # layout and index maps from the independent restatement oracle/indexing.py, no libasset_b200.so in this process);
    # tests/test_standalone_oracle.py shows the two builds agree bit for bit on pattern and input.  Otherwise (block-
    # constant control, other ODEs) the product's host code transcribes the phase; no GPU call is made either way.
    built_by = "oracle/standalone.py (independent of the product)"
    if a.ode in standalone.SIZES and a.control_mode != "BlockConstant":
        funcs, X_, dims = standalone.describe(a.ode, a.mode, nseg, control_mode=a.control_mode, extras=not a.no_extras)

        def make():
            return (standalone.instantiate(funcs, dims), X_) + standalone.multipliers(dims)
    else:
        import asset_asrl_b200  # noqa: F401
        from asset_asrl_b200 import workloads as W
        ph = W.make_phase(a.ode, a.mode, nseg, extras=not a.no_extras, control_mode=a.control_mode)
        nlp = ph.transcribe()
        built_by = "the product's host transcription, mirrored into the oracle"

        def make():
            return (oracle.Oracle.from_problem(nlp), ph.makeSolverInput()) + tuple(W.multipliers(nlp))
    threads = max(1, min(oracle.hardware_threads(), os.cpu_count() or 1))
    rates = {}
    for variant in ("ieee", "fast"):
        try:
            oracle.use_variant(variant)
            orc, X, LE, LI = make()
        except Exception as e:  # no compiler on the box: the IEEE copy travels prebuilt, the native one may be missing
            rates[variant] = None
            sys.stderr.write(f"bench.py: oracle variant {variant} unavailable: {e}\n")
            continue
        times = []
        for k in range(warmup + steps):
            t0 = time.perf_counter()
            orc.eval("KKT", 1.0, X, LE, LI, threads=threads)
            dt = time.perf_counter() - t0
            if k >= warmup:
                times.append(dt)
        rates[variant] = float(np.mean(times))
        del orc
    oracle.use_variant("ieee")
    t = min(v for v in rates.values() if v)
    return {"value": nseg / t, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{nseg} segments of {workload_name(a)} per step, {steps} timed evalKKT calls, "
                      f"{threads} threads over contiguous application chunks; oracle/oracle.cpp (reference binary unbuildable here); "
                      f"value = the faster of the two builds; problem built by {built_by}",
            "segments_per_sec_ieee_build": nseg / rates["ieee"] if rates.get("ieee") else None,
            "segments_per_sec_reference_flags_build": nseg / rates["fast"] if rates.get("fast") else None,
            "segments": nseg, "ms_per_step": t * 1e3}


def run_reference(a):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    # the driver's step counts are honoured; the workload is the GPU arm's (same segment count) as long as a step stays
    # within seconds on the host cores (200 k low-thrust LGL5 segments take about one second on 16 threads)
    steps, warmup = max(1, a.steps), max(0, a.warmup)
    full = a.segments <= 400000
    cb = cpu_eval_rate(a, steps, warmup, full_size=full)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "segments_per_gpu": cb["segments"], "segments_total": cb["segments"],
                       "eval": "evalKKT", "same_segments_as_gpu_arm": full},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "segments_per_sec_ieee_build",
                                                "segments_per_sec_reference_flags_build")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ parity leg
def parity_check(nlp, ph, nseg_total, world, rank, X, LE, LI, sc, dist=None, windowed=False):
    """Not timed: one evalKKT of the bench problem against oracle windows (oracle/windows.py) of 128 segments at both
    mesh ends, the middle and every rank cut; the ranks' disjoint slices are summed over NCCL first.  Returns the worst
    relative difference (rank 0) -- the value evidence that travels with the throughput line."""
    import torch
    from asset_asrl_b200.Solvers import EVAL_KKT
    from oracle.windows import WindowCheck
    pd = ph.dims()
    vs = pd["XtUVars"] * (pd["cardinalStates"] - 1)
    if pd["blocked"]:
        return None
    if windowed:
        # rank-local analysis: every rank checks the rows it owns on its side of its two cuts (or mesh ends) and in the
        # middle of its range, in its own CSR pattern; nothing to sum, the worst difference is maximised over the ranks
        lo, hi = nseg_total * rank // world, nseg_total * (rank + 1) // world
        windows = [(lo * vs, min(lo + 64, hi) * vs), (max((lo + hi) // 2 - 32, lo) * vs, min((lo + hi) // 2 + 32, hi) * vs),
                   (max(hi - 64, lo) * vs, hi * vs)]
        wc = WindowCheck(nlp, windows, skip_vars=nlp.window_far_vars(),
                         own_apps=lambda f: (nlp.window_info(f)["own_lo"], nlp.window_info(f)["own_hi"]))
        nlp.set_inputs(X, LE, LI)
        nlp.eval_device(EVAL_KKT, 0.7)
        nlp.sync()
        slices = {}
        for k in ("KKT", "AGX", "PGX", "FXE", "FXI"):
            idx = torch.from_numpy(wc.index[k]).cuda()
            slices[k] = (nlp.device_tensor(k)[idx] if idx.numel() else torch.zeros(0, dtype=torch.float64, device="cuda")).cpu().numpy()
        worst = wc.compare(slices, 0.7, X, LE, LI, sc)
        keys = ("KKT", "AGX", "PGX", "FXE", "FXI")
        t = torch.tensor([worst.get(k, 0.0) for k in keys], device="cuda", dtype=torch.float64)
        rows = torch.tensor([float(worst["rows"]), float(worst["applications"])], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(rows)
        if rank != 0:
            return None
        per = {k: float(v) for k, v in zip(keys, t.tolist())}
        return {"max_rel": max(per.values()), "per_array": per, "windows": 3 * world, "rows": int(rows[0].item()),
                "applications": int(rows[1].item()),
                "how": "per rank: oracle windows of 64 segments on its side of both cuts (or mesh ends) and in the middle of its "
                       "range, compared in the rank's own CSR pattern; maximum over the ranks; tolerance 1e-12 of the window's max-norm"}
    centres = [64, nseg_total // 2, nseg_total - 64] + [nseg_total * r // world for r in range(1, world)]
    windows = [(max(c - 64, 0) * vs, min(c + 64, nseg_total) * vs) for c in centres]
    wc = WindowCheck(nlp, windows)
    nlp.set_inputs(X, LE, LI)
    nlp.eval_device(EVAL_KKT, 0.7)
    nlp.sync()
    slices = {}
    for k in ("KKT", "AGX", "PGX", "FXE", "FXI"):
        idx = torch.from_numpy(wc.index[k]).cuda()
        t = nlp.device_tensor(k)[idx] if idx.numel() else torch.zeros(0, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t)
        slices[k] = t.cpu().numpy()
    if rank != 0:
        return None
    worst = wc.compare(slices, 0.7, X, LE, LI, sc)
    return {"max_rel": max(worst[k] for k in ("KKT", "AGX", "PGX", "FXE", "FXI") if k in worst), "per_array": {k: worst[k] for k in worst if k in ("KKT", "AGX", "PGX", "FXE", "FXI")},
            "windows": len(windows), "rows": worst["rows"], "applications": worst["applications"],
            "how": "oracle windows of 128 segments at both mesh ends, the middle and every rank cut; tolerance 1e-12 of the window's max-norm"}


# ------------------------------------------------------------------------------------------------ GPU arm
def _lib_cut_count(nlp):
    from asset_asrl_b200._lib import lib
    return lib.ab_plan_cut_count(nlp.plan)


def run_b200(a):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the KKT path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import asset_asrl_b200  # noqa: F401
    from asset_asrl_b200 import workloads as W
    from asset_asrl_b200 import Solvers as _S
    EVAL_KKT = getattr(_S, "EVAL_" + a.eval)  # the step's entry point (evalKKT unless --eval says otherwise)
    if a.fused_partial:
        os.environ["ASSET_B200_FUSED_PARTIAL"] = "1"
    if a.fused_blocked:
        os.environ["ASSET_B200_FUSED_BLOCKED"] = "1"
    if a.eval != "KKT":
        a.no_parity = a.no_e2e = True

    t_setup = time.time()
    sharded = world > 1 and a.shard in ("one-phase", "windows")
    windows = world > 1 and a.shard == "windows"
    requested_segments = a.segments
    if windows:
        # one phase of world x segments; every rank transcribes it, keeps its own segments plus a halo and analyses those.
        # All ranks of a node share its host memory: about 70 KB per segment at the peak of the analysis and plan build.
        try:
            import psutil
            fit = int(0.6 * psutil.virtual_memory().available / (70e3 * world)) // 25000 * 25000
        except ImportError:
            fit = a.segments
        t = torch.tensor([min(fit, a.segments)], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)  # every rank must build the same phase
        a.segments = max(25000, int(t.item()))
        ph = W.make_phase(a.ode, a.mode, a.segments * world, extras=not a.no_extras, seed=12345, control_mode=a.control_mode)
        nlp = ph.transcribe_window(rank, world)
        d = nlp.dims()
        X = ph.makeSolverInput()
        LE, LI = W.multipliers(nlp, seed=67890)
        nlp.to_device(local, rank=rank, world=world)
    elif sharded:
        # one phase of world x segments; every rank builds the same problem and evaluates its share of every function
        cap = a.max_total_segments
        try:  # every rank holds the whole analysed problem on the host: stay well inside the box's memory
            import psutil
            cap = min(cap, int(0.5 * psutil.virtual_memory().available / (45e3 * world)))
        except ImportError:
            pass
        a.segments = max(1000, min(a.segments, cap // world))
        ph = W.make_phase(a.ode, a.mode, a.segments * world, extras=not a.no_extras, seed=12345, control_mode=a.control_mode)
        nlp = ph.transcribe()
        d = nlp.dims()
        X = ph.makeSolverInput()
        LE, LI = W.multipliers(nlp, seed=67890)
        nlp.to_device(local, rank=rank, world=world)
    else:
        ph = W.make_phase(a.ode, a.mode, a.segments, extras=not a.no_extras, seed=12345 + rank, control_mode=a.control_mode)
        nlp = ph.transcribe()
        d = nlp.dims()
        X = ph.makeSolverInput()
        LE, LI = W.multipliers(nlp, seed=67890 + rank)
        nlp.to_device(local)
    sc = np.linspace(0.5, 1.5, d["n_solver"])
    nlp.set_solver_coeffs(sc)
    t_setup = time.time() - t_setup
    parity = None
    if not a.no_parity and (sharded or world == 1):
        parity = parity_check(nlp, ph, a.segments * (world if sharded else 1), world if sharded else 1, rank, X, LE, LI, sc,
                              dist if world > 1 else None, windows)
        bad = parity is not None and not parity["max_rel"] < 1e-12
        if world > 1:  # rank 0 holds the verdict; every rank must stop, or the others wait in the next collective
            flag = torch.tensor([int(bad)], device="cuda", dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            bad = int(flag.item()) == 1
        if bad:
            raise SystemExit(f"bench.py: the CUDA path differs from the oracle: {parity}")
    stream = torch.cuda.ExternalStream(nlp.stream(), device=local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput
    nlp.set_inputs(X, LE, LI)
    nlp.sync()
    for _ in range(a.warmup):
        nlp.eval_device(EVAL_KKT, 1.0)
    nlp.sync()
    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(60):  # keep the device busy until the sampler has delivered, so that its lines are taken under load
        for _ in range(50):
            nlp.eval_device(EVAL_KKT, 1.0)
        nlp.sync()
        done = len(clocks.lines) >= 3
        if world > 1:
            # every evaluation of a sharded plan contains a collective: the ranks must leave this loop together
            flag = torch.tensor([int(done)], device="cuda", dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            done = int(flag.item()) == 1
        if done:
            break
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    w0 = time.time()
    e0.record(stream)
    for _ in range(a.steps):
        nlp.eval_device(EVAL_KKT, 1.0)
    e1.record(stream)
    nlp.sync()
    barrier()
    w1 = time.time()
    ms_total = e0.elapsed_time(e1)
    launches = nlp.launches() * a.steps
    specialised = nlp.specialised_launches() * a.steps
    clk = clocks.stop(w0, w1)
    # per-launch device times: a second pass with an event pair around every launch (this serialises the small
    # launches that overlap the bulk kernels in the timed region above, so it is kept out of it)
    nlp.profile(2)
    for _ in range(max(1, min(a.steps, 5))):
        nlp.eval_device(EVAL_KKT, 1.0)
    prof = nlp.profile_times()
    nlp.profile(0)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / a.steps
    value = world * a.segments / (ms_step * 1e-3)

    # ---- per-kernel device time (CUDA events on the plan's stream, averaged over the timed region)
    agg = {}
    for label, ms in prof:
        agg.setdefault(label, []).append(ms)
    kernels = {k: float(np.mean(v)) for k, v in agg.items()}
    dom = max(kernels, key=kernels.get)
    ir, orr = None, None
    for f in range(d["nfuncs"]):
        fi = nlp.function_info(f)
        if nlp.tape_describe(fi["tape"]).split(" ", 1)[0] in ("lgl_defect", "trap_defect"):
            ir, orr = fi["ir"], fi["orows"]
    cs = W.CARDINALS[a.mode]
    n = (ir - 0) // cs
    b_seg_defect = 8 * (2 * (cs - 1) * n + 2 * orr + orr * ir + ir * (ir + 1) // 2 - n * (n + 1) // 2)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    if a.eval != "KKT":  # inputs once, the arrays the mode produces once
        prod = {"KKTNO": (1, 0, 1, 1), "SOE": (1, 0, 0, 1), "AUG": (1, 1, 1, 1), "RHS": (0, 1, 1, 1), "OGC": (0, 1, 0, 1), "OCC": (0, 0, 0, 1),
                "OBJ": (0, 0, 0, 0)}[a.eval]
        b_alg_eval = 8 * (d["nvars"] + d["neq"] + d["niq"] + prod[0] * (d["nnz"] - W._solver_only_slots(nlp)) + (prod[1] + prod[2]) * d["nvars"]
                          + prod[3] * (d["neq"] + d["niq"])) // (world if sharded else 1)
    elif windows:  # per GPU: the rank's share of the inputs and outputs, its segments' CSR rows
        b_alg_eval = 8 * ((3 * d["nvars"] + 2 * (d["neq"] + d["niq"])) // world + nlp.fused_info()["slots_per_segment"] * a.segments)
    else:
        b_alg_eval = W.algorithmic_bytes(nlp) // (world if sharded else 1)  # per GPU
    if dom == "defect":
        bytes_launch = b_seg_defect * a.segments
    else:  # another kernel dominates: charge it the whole evaluation's algorithmic bytes (upper bound on its share)
        bytes_launch = b_alg_eval
    achieved = bytes_launch / (kernels[dom] * 1e-3) / 1e9
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tj.get(workload_name(a))
        if ent and ent.get("kernel") == dom:
            traffic = ent["dram_bytes_per_segment"] * a.segments
    except (OSError, ValueError, KeyError):
        pass
    # fp64 pipe: DFMAs of the fused programs against the measured DFMA peak (tools/dfma_bench.cu) -- far from binding
    fp64 = None
    try:
        pk = json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["fp64_tflops"]
        fi = nlp.fused_info()
        if fi["ok"] and dom == "fused":
            fl = 2.0 * fi["fma"] * (nlp.fused_segments() or a.segments)
            fp64 = {"tflops": fl / (kernels[dom] * 1e-3) / 1e12, "peak_tflops": pk, "frac": fl / (kernels[dom] * 1e-3) / 1e12 / pk,
                    "dfma_per_segment": fi["fma"]}
    except (OSError, ValueError, KeyError):
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": "profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per segment from the "
                                                      "committed ncu --set full capture, scaled by the segment count (not re-measured in this run)"
                if traffic else None,
                "fp64": fp64, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_launch,
                "kernel_ms": kernels[dom], "kernel_share_of_step": kernels[dom] / ms_step,
                "eval_algorithmic_bytes": b_alg_eval, "eval_frac": b_alg_eval / (ms_step * 1e-3) / 1e9 / peak}

    # ---- end to end through the host-buffer C ABI call (pinned buffers)
    e2e = None
    if not a.no_e2e:
        def pinned(nel):
            return torch.empty(max(nel, 1), dtype=torch.float64).pin_memory().numpy()[:nel]
        hX, hLE, hLI = pinned(d["nvars"]), pinned(d["neq"]), pinned(d["niq"])
        hX[:], hLE[:], hLI[:] = X, LE, LI
        h2d = 8 * (d["nvars"] + d["neq"] + d["niq"])
        upload_note = "pinned host X/LE/LI in (replicated)"
        if sharded:  # compact pinned buffers holding this rank's output ranges only
            rg = nlp.output_ranges()
            out = [pinned(max(sum(hi - lo for (lo, hi) in rg[k]), 1)) for k in ("val", "PGX", "AGX", "FXE", "FXI", "KKT")]
            host_call = lambda *args: nlp.eval_host_sharded(*args, compact=True)  # noqa: E731
            if windows and not a.full_upload:
                # A rank of a windowed plan needs its own segments' slice of the iterate (+ halo, + far variables), not
                # the phase's: upload only the ranges its functions read.  Checked here before it is timed: the device
                # copy of the iterate is poisoned with NaN, then a sliced evaluation must reproduce the outputs of a
                # full-upload evaluation bit for bit; otherwise the run falls back to the full upload and says so.
                # Collectives (the cut-set reduce inside an evaluation) are entered by every rank or by none: each stage's
                # verdict is agreed on across the ranks before the next stage starts.
                def agreed(v):
                    flag = torch.tensor([int(v)], device="cuda", dtype=torch.int32)
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                    return int(flag.item()) == 1
                from asset_asrl_b200 import _lib as _ablib
                host_call(EVAL_KKT, 1.0, hX, hLE, hLI, *out)
                nlp.sync()
                want = [o.copy() for o in out]
                ok = 0
                try:  # stage 1, rank-local: poison the device iterate, upload the slices
                    for nm in ("X", "LE", "LI"):
                        t = nlp.device_tensor(nm)
                        if t.numel():
                            t.fill_(float("nan"))
                    torch.cuda.synchronize()
                    _ablib.check0(_ablib.lib.ab_plan_set_inputs_sliced(nlp.plan, _ablib.dptr(hX), _ablib.dptr(hLE), _ablib.dptr(hLI)))
                    nlp.sync()
                    ok = 1
                except Exception as e:  # noqa: BLE001 - any failure of the new path means: use the old one
                    upload_note = f"sliced upload failed ({type(e).__name__}: {e}); " + upload_note
                all_ok = agreed(ok)
                if all_ok:  # stage 2, collective: evaluate on what was uploaded, read back, compare
                    nlp.eval_device(EVAL_KKT, 1.0)
                    ok = 0
                    try:
                        _ablib.check0(_ablib.lib.ab_plan_get_outputs_packed(nlp.plan, *[_ablib.dptr(out[k]) for k in (5, 1, 2, 3, 4, 0)]))
                        nlp.sync()
                        ok = int(all(np.array_equal(w_, o_) for w_, o_ in zip(want, out)))
                    except Exception as e:  # noqa: BLE001
                        upload_note = f"sliced upload check failed ({type(e).__name__}: {e}); " + upload_note
                    all_ok = agreed(ok)
                del want
                flag_all = all_ok
                if flag_all:
                    host_call = lambda *args: nlp.eval_host_sharded(*args, compact=True, sliced=True)  # noqa: E731
                    ir_ = nlp.input_ranges()
                    h2d = 8 * sum(hi - lo for rr in ir_.values() for (lo, hi) in rr)
                    upload_note = ("pinned host X/LE/LI, of which only the ranges this rank's functions read are copied "
                                   f"({sum(len(rr) for rr in ir_.values())} copies; checked against a full upload with the device iterate poisoned)")
                elif ok == 1:
                    upload_note = "sliced upload disagreed with the full upload on another rank; " + upload_note
                elif "failed" not in upload_note:
                    upload_note = "sliced upload did not reproduce the full upload's outputs; " + upload_note
        else:
            out = [pinned(1), pinned(d["nvars"]), pinned(d["nvars"]), pinned(d["neq"]), pinned(d["niq"]), pinned(d["nnz"])]
            host_call = nlp.eval_host
        ksteps = max(1, min(a.steps, 5))
        for _ in range(2):
            host_call(EVAL_KKT, 1.0, hX, hLE, hLI, *out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(ksteps):
            host_call(EVAL_KKT, 1.0, hX, hLE, hLI, *out)
        barrier()
        dt = (time.perf_counter() - t0) / ksteps
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        d2h = 8 * (1 + 2 * d["nvars"] + d["neq"] + d["niq"] + d["nnz"])
        note = "ab_plan_eval_host: pinned host X/LE/LI in, val/PGX/AGX/FXE/FXI/KKT values out"
        if sharded:
            d2h = 8 * sum(hi - lo for rr in nlp.output_ranges().values() for (lo, hi) in rr)
            note = ("per rank: " + upload_note + ", evaluation of the rank's share, NCCL cut-set reduce, "
                    "the rank's output ranges out; bytes are rank 0's")
        e2e = {"value": world * a.segments / dt, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3, "steps": ksteps, "note": note}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        cb = cpu_eval_rate(a, 3, 1)
        cpu = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "segments_per_sec_ieee_build",
                                  "segments_per_sec_reference_flags_build")}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak" if a.segments == requested_segments else "neither (segments per GPU were reduced to fit the host analysis)",
                "segments_per_gpu_reduced": a.segments != requested_segments, "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": workload_name(a), "segments_per_gpu": a.segments, "segments_total": world * a.segments,
                           "eval": "eval" + a.eval, "nvars": d["nvars"], "neq": d["neq"], "niq": d["niq"], "nnz": d["nnz"],
                           "kkt_evals_per_sec": 1e3 / ms_step,
                           "sharding": (("one phase, rank-local analysis (own segments + halo of 8), each rank fills the CSR rows of "
                                         "its segments, NCCL all-reduce of the cut set (" if windows else
                                         "one phase split into contiguous application ranges, every rank analyses the whole problem, "
                                         "NCCL all-reduce of the cut set (") + str(int(_lib_cut_count(nlp))) + " doubles)") if sharded else
                                       ("one independent phase per rank, no collective" if world > 1 else "single GPU"),
                           "l2": f"no flush: per-eval working set {b_alg_eval / 1e6:.0f} MB of values (+ index maps) > 126 MB L2",
                           "setup_s": round(t_setup, 2)},
                "parity_max_rel": parity["max_rel"] if parity else None, "parity": parity,
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "specialised_launches": specialised, "clocks": clk,
                "kernels_ms": kernels}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
