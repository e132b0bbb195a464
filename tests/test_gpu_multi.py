"""Sharded evaluation on the GPU against the CPU oracle.

* one device, two plans (rank 0 and rank 1 of a world of 2; the cut buffers are added by hand) -- runs on the 1-GPU box;
* two devices, two processes, NCCL all-reduce of the cut buffer -- skipped when fewer than two GPUs are visible."""
import os
import socket
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import oracle  # noqa: E402
from helpers import phase_zoo, problem_inputs, rel_err, TOL  # noqa: E402

from asset_asrl_b200 import workloads as W  # noqa: E402
from asset_asrl_b200.Solvers import EVAL_KKT, EVAL_RHS, NonLinearProgram  # noqa: E402

pytestmark = pytest.mark.gpu
ZOO = phase_zoo()
KEYS = ("PGX", "AGX", "FXE", "FXI", "KKT")


def build(name):
    ph = ZOO[name]() if name in ZOO else W.make_phase("lowthrust", "LGL5", int(name))
    nlp = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp)
    return ph, nlp, X, LE, LI, sc


def clone_problem(ph):
    return ph.transcribe()


@pytest.mark.parametrize("name", ["lowthrust_LGL5", "regions_LGL7", "regions_LGL5_block_constant",
                                  "ocp_multispacecraft_LGL5_block_constant", "300"])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_plans_on_one_device(name, world):
    import torch
    ph, nlp0, X, LE, LI, sc = build(name)
    ref = oracle.Oracle.from_problem(nlp0).eval("KKT", 0.7, X, LE, LI, sc)
    plans = []
    for r in range(world):
        nlp = clone_problem(ph)
        nlp.to_device(0, rank=r, world=world)
        nlp.set_solver_coeffs(sc)
        nlp.set_inputs(X, LE, LI)
        plans.append(nlp)
    for mode, keys in ((EVAL_KKT, KEYS + ("val",)), (EVAL_RHS, ("PGX", "AGX", "FXE", "FXI", "val"))):
        for p in plans:
            _ = p.world
            p.world = 1  # no NCCL in this test: the reduction is done by hand below
            p.eval_device(mode, 0.7)
            p.world = world
            p.cut_pack()
            p.sync()
        total = sum(p.cut_tensor().clone() for p in plans)
        for p in plans:
            p.cut_tensor().copy_(total)
            torch.cuda.synchronize()
            p.cut_unpack()
            p.sync()
        outs = [p.get_outputs() for p in plans]
        for k in keys:
            if k == "val":
                tot = sum(o["val"] for o in outs)
                assert abs(tot - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0)
                continue
            stack = np.stack([o[k] for o in outs])
            assert np.all((stack != 0).sum(axis=0) <= 1), (k, "an output is held by more than one rank")
            assert rel_err(stack.sum(axis=0), ref[k]) < TOL, (name, mode, k)
            if name == "300" and mode == EVAL_KKT and k == "KKT":
                assert sum(p.launches() for p in plans) > 0
                nz = [(o["KKT"] != 0).sum() for o in outs]
                assert min(nz) > 0.2 * max(nz), "the value array is not split evenly"



def test_world8_sharded_plans_at_20k_segments():
    """World 8 (the north-star box): eight sharded plans of one 20 000-segment phase on one device, the cut set reduced by
    hand, every output against the full oracle and -- through oracle windows -- around each of the seven cuts."""
    import torch
    from oracle.windows import check_windows
    world, nseg = 8, 20000
    ph = W.make_phase("lowthrust", "LGL5", nseg)
    nlp0 = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(nlp0)
    ref = oracle.Oracle.from_problem(nlp0).eval("KKT", 0.7, X, LE, LI, sc, threads=oracle.hardware_threads())
    plans = []
    for r in range(world):
        nlp = clone_problem(ph)
        nlp.to_device(0, rank=r, world=world)
        nlp.set_solver_coeffs(sc)
        nlp.set_inputs(X, LE, LI)
        plans.append(nlp)
    for p in plans:
        p.world = 1
        p.eval_device(EVAL_KKT, 0.7)
        p.world = world
        p.cut_pack()
        p.sync()
    total = sum(p.cut_tensor().clone() for p in plans)
    for p in plans:
        p.cut_tensor().copy_(total)
        torch.cuda.synchronize()
        p.cut_unpack()
        p.sync()
    outs = [p.get_outputs() for p in plans]
    summed = {}
    for k in KEYS:
        stack = np.stack([o[k] for o in outs])
        assert np.all((stack != 0).sum(axis=0) <= 1), (k, "an output is held by more than one rank")
        summed[k] = stack.sum(axis=0)
        assert rel_err(summed[k], ref[k]) < TOL, k
    tot = sum(o["val"] for o in outs)
    assert abs(tot - ref["val"]) <= TOL * max(abs(ref["val"]), 1.0)
    assert all(p.fused_segments() > 0 for p in plans)
    windows = [(max(nseg * r // world - 64, 0) * 20, (nseg * r // world + 64) * 20) for r in range(1, world)]
    worst = check_windows(nlp0, summed, windows, 0.7, X, LE, LI, sc)
    for k in KEYS:
        assert worst[k] < TOL, (k, worst)

def _nccl_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        ph = W.make_phase("lowthrust", "LGL5", 2000)
        nlp = ph.transcribe()
        X = ph.makeSolverInput()
        LE, LI, sc = problem_inputs(nlp)
        nlp.to_device(rank, rank=rank, world=world)
        nlp.set_solver_coeffs(sc)
        out = nlp.evalKKT(0.7, X, LE, LI)
        worst = 0.0
        ref = oracle.Oracle.from_problem(nlp).eval("KKT", 0.7, X, LE, LI, sc) if rank == 0 else None
        for k in KEYS + ("val",):
            t = torch.from_numpy(np.atleast_1d(np.asarray(out[k], dtype=np.float64)).copy()).cuda()
            dist.all_reduce(t)
            if rank == 0:
                worst = max(worst, rel_err(t.cpu().numpy(), np.atleast_1d(ref[k])))
        if rank == 0:
            q.put(worst)
    finally:
        dist.destroy_process_group()


def test_two_gpus_nccl():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    assert q.get() < TOL
