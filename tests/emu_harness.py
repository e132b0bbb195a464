"""Runs the fused kernels' PTX on the CPU emulator (ptx_emu.py) for a few threads and compares what they write with the
oracle.  Test infrastructure."""
import numpy as np

from ptx_emu import Kernel


def stage_solver_offset(nlp):
    """First staging slot of the solver-owned coefficients (csrc/nlp.cpp: they follow the staged user triplets)."""
    d = nlp.dims()
    lists, _ = nlp.gather_lists()
    loc = nlp.kkt_locations()
    t0 = int(loc[d["n_user"]])
    best = -1
    for ptr, src, tgt in lists["KKT"]:
        hit = np.nonzero(tgt == t0)[0]
        for h in hit:
            best = max(best, int(src[ptr[h]:ptr[h + 1]].max()))
    assert best >= 0
    return best


def emulate_fused(nlp, X, LE, LI, sc, obj_scale, rels, residual_mode=None):
    """Returns {array name: (values, written mask)} after running every fused piece for the threads in `rels`."""
    d = nlp.dims()
    info = nlp.fused_info()
    assert info["ok"], info["why"]
    _, stage_size = nlp.gather_lists()
    stage = np.zeros(max(stage_size, 1))
    s0 = stage_solver_offset(nlp)
    stage[s0:s0 + d["n_solver"]] = sc
    appc = np.concatenate([nlp.function_appconst(f).ravel() for f in range(d["nfuncs"])] + [np.zeros(1)])
    nan = np.nan
    out = {"KKT": np.full(d["nnz"], nan), "FXE": np.full(d["neq"], nan), "FXI": np.full(max(d["niq"], 1), nan),
           "PGX": np.full(d["nvars"], nan), "AGX": np.full(d["nvars"], nan)}
    arrays = {0: np.asarray(X, dtype=np.float64), 1: np.asarray(LE, dtype=np.float64), 2: out["KKT"], 3: out["FXE"], 4: stage,
              5: appc, 6: np.asarray(LI, dtype=np.float64) if d["niq"] else np.zeros(1), 7: out["FXI"], 8: out["PGX"], 9: out["AGX"]}
    if residual_mode is None:
        ptxs = [nlp.fused_ptx(p) for p in range(int(info["pieces"]))]
    elif residual_mode >= 4:
        ptxs = [nlp.fused_residual_ptx(residual_mode)]
    else:  # partial-matrix modes: one kernel per row block, like evalKKT
        ptxs = []
        while True:
            t = nlp.fused_mode_ptx(residual_mode, len(ptxs))
            if t is None:
                break
            ptxs.append(t)
        assert len(ptxs) == int(info["pieces"])
    for ptx in ptxs:
        k = Kernel(ptx)
        for rel in rels:
            k.run(arrays, rel, obj_scale)
    res = {k: (v, ~np.isnan(v)) for k, v in out.items()}
    res["stage"] = stage
    return res
