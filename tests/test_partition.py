"""Multi-GPU partition of one problem: host logic on CPU, world_size 2 over gloo.

Each rank derives its share (application ranges, restricted gather lists, cut set) independently; the test emulates
the device side in numpy -- a rank's staging buffer holds values only in the slots its own applications write -- runs
the same pack / all-reduce / unpack sequence the GPU path runs over NCCL, and checks that the per-rank outputs are
disjoint and sum to the unpartitioned fixed-order gather."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import phase_zoo  # noqa: E402

from asset_asrl_b200.Solvers import OUT_ARRAYS  # noqa: E402

ZOO = phase_zoo()
CASES = ["lowthrust_LGL5", "regions_LGL7", "regions_LGL5_block_constant", "ocp_multispacecraft_LGL5_block_constant"]


def gather(lists, stage, size):
    out = np.zeros(size)
    written = np.zeros(size, dtype=bool)
    for (ptr, src, tgt) in lists:
        for t in range(len(tgt)):
            s = 0.0
            for k in range(ptr[t], ptr[t + 1]):
                s += stage[src[k]]
            out[tgt[t]] = s
            written[tgt[t]] = True
    return out, written


def out_sizes(d):
    return {"KKT": d["nnz"], "PGX": d["nvars"], "AGX": d["nvars"], "FXE": d["neq"], "FXI": d["niq"], "val": 1}


def source_ranks(nlp, world):
    """rank writing each staging slot, from the union of the ranks' own gather lists (-1: nobody's, e.g. unused)."""
    _, stage = nlp.gather_lists()
    owner = np.full(stage, -1)
    for r in range(world):
        part = nlp.partition(r, world)
        for name in OUT_ARRAYS:
            for (ptr, src, tgt) in part["gathers"][name]:
                assert np.all(owner[src] == -1), "a staging slot is summed by two ranks"
                owner[src] = r
    return owner


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("world", [2, 3])
def test_partition_covers_every_source_once(name, world):
    nlp = ZOO[name]().transcribe()
    d = nlp.dims()
    full, stage = nlp.gather_lists()
    owner = source_ranks(nlp, world)
    used = np.zeros(stage, dtype=bool)
    for nm in OUT_ARRAYS:
        for (ptr, src, tgt) in full[nm]:
            used[src] = True
    assert np.array_equal(used, owner >= 0)
    # application ranges tile every function
    rngs = [nlp.partition(r, world)["app_ranges"] for r in range(world)]
    for f in range(d["nfuncs"]):
        assert rngs[0][f, 0] == 0 and rngs[-1][f, 1] == nlp.function_info(f)["nappl"]
        for r in range(1, world):
            assert rngs[r][f, 0] == rngs[r - 1][f, 1]
    # the cut set is identical on every rank and its owner contributes
    parts = [nlp.partition(r, world) for r in range(world)]
    for nm in OUT_ARRAYS:
        t0, o0, _ = parts[0]["cuts"][nm]
        for r in range(world):
            t, o, m = parts[r]["cuts"][nm]
            assert np.array_equal(t, t0) and np.array_equal(o, o0)
            assert np.all(m[o == r] == 1)


def _worker(rank, world, port, name, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nlp = ZOO[name]().transcribe()
        d = nlp.dims()
        sizes = out_sizes(d)
        full, stage_size = nlp.gather_lists()
        rng = np.random.default_rng(99)
        stage_full = rng.uniform(-1, 1, stage_size)
        owner = source_ranks(nlp, world)
        stage = np.where(owner == rank, stage_full, 0.0)  # what this rank's kernels would have written
        part = nlp.partition(rank, world)
        outs = {}
        for nm in OUT_ARRAYS:
            outs[nm], _ = gather(part["gathers"][nm], stage, sizes[nm])
        # pack / all-reduce / unpack, exactly the device sequence (Plan::cut_pack, NCCL, Plan::cut_unpack)
        buf = np.concatenate([np.where(part["cuts"][nm][2] == 1, outs[nm][part["cuts"][nm][0]], 0.0) for nm in OUT_ARRAYS])
        tbuf = torch.from_numpy(buf.copy())
        dist.all_reduce(tbuf, op=dist.ReduceOp.SUM)
        buf = tbuf.numpy()
        off = 0
        for nm in OUT_ARRAYS:
            t, o, m = part["cuts"][nm]
            seg = buf[off:off + len(t)]
            off += len(t)
            outs[nm][t[o == rank]] = seg[o == rank]
            outs[nm][t[(o != rank) & (m == 1)]] = 0.0
        # every output non-zero on one rank only; the sum over ranks is the unpartitioned result
        tot = {}
        for nm in OUT_ARRAYS:
            nzc = torch.from_numpy((outs[nm] != 0).astype(np.int64))
            dist.all_reduce(nzc)
            assert int(nzc.max()) <= 1, (nm, "an output is held by two ranks")
            s = torch.from_numpy(outs[nm].copy())
            dist.all_reduce(s)
            tot[nm] = s.numpy()
        if rank == 0:
            worst = 0.0
            for nm in OUT_ARRAYS:
                ref, _ = gather(full[nm], stage_full, sizes[nm])
                worst = max(worst, float(np.abs(tot[nm] - ref).max()) if ref.size else 0.0)
            q.put(worst)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["lowthrust_LGL5", "regions_LGL5_block_constant"])
def test_world2_gloo_cut_reduce(name):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get() < 1e-13  # cut targets are summed as (partial_0 + partial_1): at most one rounding apart from the fixed order
