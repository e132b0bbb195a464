"""Index maps of the product against the independent restatement of the reference's PhaseIndexer and transcription order
(tests/ref_indexing.py): Vindex / Cindex of every function of every phase_zoo entry, bit-exact, plus problem sizes."""
import numpy as np
import pytest

import asset_asrl_b200  # noqa: F401
from asset_asrl_b200 import OptimalControl as oc
from asset_asrl_b200 import workloads as W

import ref_indexing as R
from helpers import phase_zoo

ZOO = phase_zoo()


def product_lists(nlp):
    """Product functions grouped as the reference's three lists, in emission order."""
    d = nlp.dims()
    lists = {0: [], 1: [], 2: []}
    for f in range(d["nfuncs"]):
        fi = nlp.function_info(f)
        V, Cc = nlp.function_indices(f)
        lists[fi["role"]].append((fi, V, Cc))
    return lists


def expand_reference(ref_list):
    """The product batches the N-1 single-column SingleMeshSpacing functions of addPartitionedEquality into one
    function with one application per column: merge consecutive reference entries of that name the same way."""
    out = []
    for fn in ref_list:
        if fn["name"] == "single_mesh_spacing" and out and out[-1]["name"] == "single_mesh_spacing":
            out[-1] = {"name": fn["name"], "Vindex": np.vstack([out[-1]["Vindex"], fn["Vindex"]]),
                       "Cindex": np.vstack([out[-1]["Cindex"], fn["Cindex"]]), "unique_constraints": True}
        else:
            out.append(dict(fn))
    return out


@pytest.mark.parametrize("name", sorted(ZOO))
def test_index_maps_match_the_reference_restatement(name):
    obj = ZOO[name]()
    nlp = obj.transcribe()
    if isinstance(obj, oc.OptimalControlProblem):
        ref, sizes = R.transcribe_ocp(obj)
    else:
        ref, sizes = R.transcribe_single_phase(obj)
    d = nlp.dims()
    assert (d["nvars"], d["neq"], d["niq"]) == (sizes["nvars"], sizes["neq"], sizes["niq"])
    got = product_lists(nlp)
    for role, rl in ((0, ref.obj), (1, ref.eq), (2, ref.iq)):
        rl = expand_reference(rl)
        assert len(got[role]) == len(rl), (role, [g[0] for g in got[role]], [r["name"] for r in rl])
        for k, ((fi, V, Cc), r) in enumerate(zip(got[role], rl)):
            assert V.shape == r["Vindex"].shape, (role, k, r["name"], V.shape, r["Vindex"].shape)
            assert np.array_equal(V, r["Vindex"]), (role, k, r["name"])
            if role != 0:
                assert np.array_equal(Cc, r["Cindex"]), (role, k, r["name"])
                assert bool(fi["unique_constraints"]) == r["unique_constraints"], (role, k, r["name"])


@pytest.mark.parametrize("blocked", [False, True])
@pytest.mark.parametrize("mode", ["LGL3", "LGL5", "LGL7", "Trapezoidal"])
def test_every_region_flag_through_the_c_abi(mode, blocked):
    """ab_phase_region_vindex (the product's make_Vindex_Cindex) against the restatement for all 14 usable flags, with
    state-only, control-only and mixed variable lists, ODE and static parameters."""
    from helpers import vdp_param_ode
    ode = vdp_param_ode()  # XV 2, UV 1, PV 1
    nseg = 5
    S = (W.CARDINALS[mode] - 1) * nseg + 1
    t = np.linspace(0.0, 1.0, S)
    T = np.stack([np.cos(t), np.sin(t), t, 0.1 * t, np.full(S, 0.5)], axis=1)
    ph = ode.phase(mode)
    if blocked:
        ph.setControlMode("BlockConstant")
    ph.setStaticParams([0.1, 0.2, 0.3])
    ph.setTraj(T, nseg)
    ix = R.make_indexer(ph)
    ix.begin_indexing(0, 0, 0)
    ocp = oc.OptimalControlProblem()
    ocp.addPhase(ph)
    flags = [R.Front, R.Back, R.FrontandBack, R.BackandFront, R.Path, R.InnerPath, R.NodalPath, R.PairWisePath,
             R.FrontNodalBackPath, R.DefectPath, R.DefectPairWisePath, R.Params, R.ODEParams, R.StaticParams]
    for reg in flags:
        for x in ([0, 2], [3], [1, 3], [0, 1, 2, 3]):
            for op, sp in (([], []), ([0], [2, 0])):
                if reg in (R.Params, R.ODEParams, R.StaticParams):
                    x_ = []
                else:
                    x_ = x
                if reg == R.FrontNodalBackPath and blocked and all(v >= 3 for v in x_) and (op or sp):
                    continue  # out-of-bounds column write in the reference (PhaseIndexer.cpp:334-335); unreachable from its callers
                want, _, _ = ix.vindex_cindex(reg, x_, op, sp, 1, 0)
                got = ocp._region_vindex(0, reg, x_, op, sp)
                assert np.array_equal(np.asarray(want), got), (reg, x_, op, sp)
