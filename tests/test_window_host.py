"""Rank-local analysis on the host (Problem::window): ownership tiles the applications, and inside the rows a rank owns
the windowed problem's CSR pattern is the whole problem's, column for column."""
import numpy as np
import pytest

from asset_asrl_b200 import workloads as W


@pytest.mark.parametrize("ode,mode,nseg,world", [("lowthrust", "LGL5", 400, 4), ("vanderpol", "LGL3", 300, 3),
                                                 ("lowthrust", "LGL7", 200, 2), ("hypersens", "Trapezoidal", 500, 5)])
def test_owned_rows_have_the_whole_problems_pattern(ode, mode, nseg, world):
    ph = W.make_phase(ode, mode, nseg)
    whole = ph.transcribe()
    d = whole.dims()
    gip, gix = whole.pattern()
    pd = ph.dims()
    vs = pd["XtUVars"] * (pd["cardinalStates"] - 1)
    owned_apps = [np.zeros(whole.function_info(f)["nappl"], dtype=np.int32) for f in range(d["nfuncs"])]
    for r in range(world):
        nlp = ph.transcribe_window(r, world)
        dl = nlp.dims()
        assert (dl["nvars"], dl["neq"], dl["niq"], dl["kktdim"]) == (d["nvars"], d["neq"], d["niq"], d["kktdim"])
        assert dl["n_user"] < 1.3 * d["n_user"] / world + 20 * d["n_user"] / nseg
        lip, lix = nlp.pattern()
        lo, hi = nseg * r // world, nseg * (r + 1) // world
        v_hi = hi * vs if r < world - 1 else pd["numStates"] * pd["XtUVars"]
        far = set(int(v) for v in nlp.window_far_vars())
        for v in range(lo * vs, v_hi):
            a0, a1, b0, b1 = int(gip[v]), int(gip[v + 1]), int(lip[v]), int(lip[v + 1])
            if v in far:  # a row every rank adds entries to (its own applications' columns): a subset of the whole row
                assert np.all(np.isin(lix[b0:b1], gix[a0:a1])), (r, v)
                continue
            assert a1 - a0 == b1 - b0 and np.array_equal(gix[a0:a1], lix[b0:b1]), (r, v)
        for f in range(d["nfuncs"]):
            wi = nlp.window_info(f)
            assert wi["nappl_whole"] == owned_apps[f].size
            owned_apps[f][wi["app_base"] + wi["own_lo"]:wi["app_base"] + wi["own_hi"]] += 1
            # index maps of the kept applications are the whole problem's
            Vw, Cw = whole.function_indices(f)
            Vl, Cl = nlp.function_indices(f)
            assert np.array_equal(Vl, Vw[wi["app_base"]:wi["app_base"] + Vl.shape[0]])
            assert np.array_equal(Cl, Cw[wi["app_base"]:wi["app_base"] + Cl.shape[0]])
        fi = nlp.fused_info()
        assert fi["ok"] == 1, fi["why"]
    for f, cnt in enumerate(owned_apps):
        assert np.all(cnt == 1), (f, "every application is owned by exactly one rank")


def test_bench_parity_windows_on_a_windowed_problem():
    """bench.py's per-rank parity leg (oracle windows on a rank-local problem), with the whole-problem oracle's values laid
    out in the rank's own CSR pattern standing in for the device output."""
    import oracle
    from oracle.windows import WindowCheck
    from helpers import problem_inputs
    nseg, world, vs = 600, 3, 20
    ph = W.make_phase("lowthrust", "LGL5", nseg)
    whole = ph.transcribe()
    X = ph.makeSolverInput()
    LE, LI, sc = problem_inputs(whole)
    ref = oracle.Oracle.from_problem(whole).eval("KKT", 0.7, X, LE, LI, sc)
    gip, gix = whole.pattern()
    for r in range(world):
        nlp = ph.transcribe_window(r, world)
        lip, lix = nlp.pattern()
        lo, hi = nseg * r // world, nseg * (r + 1) // world
        far = set(int(v) for v in nlp.window_far_vars())
        kkt = np.zeros(nlp.dims()["nnz"])
        for v in range(lo * vs, hi * vs):
            if v in far:
                continue
            kkt[lip[v]:lip[v + 1]] = ref["KKT"][gip[v]:gip[v + 1]]
        out = {"KKT": kkt, "AGX": ref["AGX"], "PGX": ref["PGX"], "FXE": ref["FXE"], "FXI": ref["FXI"]}
        windows = [(lo * vs, min(lo + 64, hi) * vs), (max((lo + hi) // 2 - 32, lo) * vs, min((lo + hi) // 2 + 32, hi) * vs),
                   (max(hi - 64, lo) * vs, hi * vs)]
        wc = WindowCheck(nlp, windows, skip_vars=nlp.window_far_vars(),
                         own_apps=lambda f: (nlp.window_info(f)["own_lo"], nlp.window_info(f)["own_hi"]))
        worst = wc.compare(wc.extract(out), 0.7, X, LE, LI, sc)
        for k in ("KKT", "AGX", "PGX", "FXE", "FXI"):
            assert worst[k] < 1e-13, (r, k, worst)


@pytest.mark.parametrize("ode,mode,nseg,world", [("lowthrust", "LGL5", 4000, 8), ("vanderpol", "LGL3", 3000, 3)])
def test_input_ranges_cover_what_a_rank_reads_and_follow_its_share(ode, mode, nseg, world):
    """ab_problem_input_ranges: the slice of X / LE / LI a rank has to receive each step (sliced upload of
    eval_host_sharded).  Everything for the whole problem; for a window every index a kept application reads lies in a
    range, ranges are ascending and disjoint, and their total is the rank's share plus halo -- not the phase."""
    ph = W.make_phase(ode, mode, nseg)
    whole = ph.transcribe()
    d = whole.dims()
    full = {"X": d["nvars"], "LE": d["neq"], "LI": d["niq"]}
    assert whole.input_ranges() == {k: ([(0, n)] if n else []) for k, n in full.items()}
    for r in range(world):
        nlp = ph.transcribe_window(r, world)
        rg = nlp.input_ranges()
        used = {"X": set(), "LE": set(), "LI": set()}
        for f in range(nlp.dims()["nfuncs"]):
            fi = nlp.function_info(f)
            V, Cc = nlp.function_indices(f)
            used["X"].update(int(v) for v in np.unique(V))
            if fi["role"]:
                used["LE" if fi["role"] == 1 else "LI"].update(int(c) for c in np.unique(Cc))
        for k, ranges in rg.items():
            for (a0, a1), (b0, b1) in zip(ranges[:-1], ranges[1:]):
                assert a0 < a1 <= b0 < b1
            cover = np.zeros(full[k] + 1, dtype=bool)
            for lo, hi in ranges:
                assert 0 <= lo < hi <= full[k]
                cover[lo:hi] = True
            idx = np.fromiter(used[k], dtype=np.int64, count=len(used[k]))
            assert cover[idx].all(), (r, k, "an index a kept application reads is not uploaded")
            total = int(cover.sum())
            if full[k]:
                assert total <= 1.25 * full[k] / world + 40 * full[k] / nseg, (r, k, total, full[k])
                assert len(ranges) <= 8, (r, k, ranges)
