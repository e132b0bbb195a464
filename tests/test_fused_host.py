"""Host side of the fused owner-writes evalKKT path (asset_asrl_b200/csrc/fused.cpp): template discovery on the analysed
problem, the cut into row-block pieces and the PTX the specialiser emits for them.  No GPU: the PTX is assembled for
sm_100a in-process.  Values are checked on the device in tests/test_gpu_fused.py.

The reference has no counterpart of this structure -- it loops over applications and adds into the shared value array
under a mutex (ConstraintFunction.h:80-109, NonLinearProgram.cpp:486-519); what has to agree is the result.
"""
import re

import pytest

from asset_asrl_b200 import workloads as W
from asset_asrl_b200.Solvers import NonLinearProgram


CASES = [
    # ode, transcription, segments, pieces the automatic rule picks
    ("lowthrust", "LGL5", 96, 2),
    ("lowthrust", "LGL3", 96, 1),
    ("lowthrust", "Trapezoidal", 96, 1),
    ("vanderpol", "LGL5", 96, 1),
    ("hypersens", "LGL5", 96, 1),
]


@pytest.mark.parametrize("ode,mode,nseg,pieces", CASES)
def test_template_and_pieces(ode, mode, nseg, pieces):
    nlp = W.make_phase(ode, mode, nseg).transcribe()
    fi = nlp.fused_info()
    assert fi["ok"] == 1, fi["why"]
    # the periodic interior excludes a few segments at either end of the mesh and nothing else
    assert 0 < fi["t_lo"] <= 8 and nseg - 8 <= fi["t_hi"] < nseg
    assert fi["pieces"] == pieces
    # every CSR slot of the owned rows has a source, so a thread's outputs are at least its region
    assert fi["outputs"] >= fi["slots_per_segment"] > 0


def test_lgl7_pieces_respect_the_parked_slot_limit():
    """The pieces are chosen by on-chip state: no piece of more than one row parks more than the limit (256 slots)."""
    nlp = W.make_phase("lowthrust", "LGL7", 64).transcribe()
    fi = nlp.fused_info()
    assert fi["ok"] == 1, fi["why"]
    assert 4 <= fi["pieces"] <= 8
    for pc in range(fi["pieces"]):
        ptx = nlp.fused_ptx(pc)
        offs = [int(m) for m in re.findall(r"\[%rsb\+(\d+)\]", ptx)]
        slots = max(offs) // 256 + 1 if offs else 0
        assert slots <= 300, (pc, slots)


def _stores(ptx):
    """Outputs a thread stores: tile-walk kernels park each in the transposition tile; direct-store kernels write runs
    with one predicated store sequence per alignment phase of the thread -- count the elements of phase 0."""
    n = len(re.findall(r"^st\.shared\.f64 \[%rso\+", ptx, flags=re.M))
    for pred, vec in re.findall(r"^(@%pm\d_\d )?st\.global(?:\.cs)?(\.v[24])?\.f64 \[%r[gh]", ptx, flags=re.M):
        if pred == "" or pred.endswith("_0 "):
            n += int(vec[2]) if vec else 1
    return n


@pytest.mark.parametrize("ode,mode", [("lowthrust", "LGL5"), ("vanderpol", "LGL3")])
def test_piece_ptx_assembles(ode, mode, monkeypatch):
    nlp = W.make_phase(ode, mode, 96).transcribe()
    fi = nlp.fused_info()
    total = 0
    for pc in range(fi["pieces"]):
        ptx = nlp.fused_ptx(pc)
        assert ".visible .entry ab_fused_kkt%d(" % pc in ptx
        info = NonLinearProgram.assemble_ptx(ptx)
        assert info["cubin_bytes"] > 0 and info["registers"] <= 255
        total += _stores(ptx)
    # the pieces together store every output of the thread once; only structural zeros outside the owned region (their
    # destinations are zero-filled when the plan is built) may be left out
    assert fi["slots_per_segment"] <= total <= fi["outputs"]
    assert fi["outputs"] - total <= 8


def test_direct_and_tile_walk_variants_store_the_same_outputs(monkeypatch):
    nlp = W.make_phase("lowthrust", "LGL5", 96).transcribe()
    fi = nlp.fused_info()
    direct = [nlp.fused_ptx(pc) for pc in range(fi["pieces"])]
    monkeypatch.setenv("ASSET_B200_FUSED_DIRECT", "0")
    walk = [nlp.fused_ptx(pc) for pc in range(fi["pieces"])]
    for a, b in zip(direct, walk):
        assert "st.global.v4.f64" in a and "st.global.v4.f64" not in b
        assert _stores(a) == _stores(b)
        # every alignment phase of a run stores the same number of elements
        for r in ("2", "4"):
            per = [sum(int(v[2]) if v else 1 for v in re.findall(r"^@%%pm%s_%d st\.global(\.v[24])?\.f64 " % (r, k), a, flags=re.M))
                   for k in range(int(r))]
            assert len(set(per)) == 1, per

