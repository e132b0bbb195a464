"""Oracle windows: value parity at full problem size without analysing the full problem on the CPU (test infrastructure;
imported only by tests/ and by bench.py's parity leg).

For a set of variable ranges ("windows") the oracle evaluates a *sub-problem* that has the full problem's dimensions
and indexing but only those function applications that touch a window variable.  Every application that contributes to
a window variable's KKT row, gradient entry or (for its own rows) constraint value is included, so inside the windows
the sub-problem's results are the full problem's results: rows of the upper-triangular KKT matrix (compared through the
two CSR patterns, which must agree column for column), AGX, PGX, and the FXE / FXI rows of the included applications.
The objective value and rows shared by applications outside the windows (accumulations, parameters touched by every
segment) are not comparable this way and are skipped; tests cover them at sizes the full oracle handles.

Two stages, so that a sharded run can sum the ranks' slices before comparing: ``WindowCheck.index`` says which
elements of each output array are looked at, ``compare`` takes exactly those elements.
"""
import numpy as np

from .binding import Oracle, hardware_threads

ARRAYS = ("KKT", "AGX", "PGX", "FXE", "FXI")


class WindowCheck:
    def __init__(self, nlp, windows, pattern=None, skip_vars=(), own_apps=None):
        """``skip_vars``: variables whose rows are not compared (a windowed problem's far variables: every rank adds to
        them).  ``own_apps(f) -> (lo, hi)``: applications of function f whose constraint rows this rank holds (a windowed
        problem keeps applications of its neighbours too; their rows live on the neighbour).

        The sub-problem is *compacted*: the variables and the equality / inequality rows its applications touch are
        renumbered 0, 1, 2, ... in ascending order (a monotone map keeps every row's column order, so CSR rows still
        compare entry by entry), which makes the oracle's cost follow the windows, not the size of the phase."""
        d = nlp.dims()
        self.windows = [(int(a), int(b)) for a, b in windows]
        lo = np.array([w[0] for w in self.windows], dtype=np.int64)
        hi = np.array([w[1] for w in self.windows], dtype=np.int64)
        picked = []
        for f in range(d["nfuncs"]):
            fi = nlp.function_info(f)
            V, Cc = nlp.function_indices(f)
            hit = np.zeros(V.shape[0], dtype=bool)
            for a, b in zip(lo, hi):
                hit |= ((V >= a) & (V < b)).any(axis=1)
            if not hit.any():
                continue
            mine = hit.copy()
            if own_apps is not None:
                lo_a, hi_a = own_apps(f)
                mine[:lo_a] = False
                mine[hi_a:] = False
            ac = nlp.function_appconst(f)
            picked.append((fi, nlp.tape_describe(fi["tape"]), V[hit].astype(np.int64), Cc[hit].astype(np.int64) if fi["role"] else None,
                           ac[hit] if ac.size else None, Cc[mine].astype(np.int64) if fi["role"] else None))
        # monotone renumbering of what the sub-problem touches
        self.vmap = np.unique(np.concatenate([p[2].ravel() for p in picked] + [np.zeros(0, dtype=np.int64)]))
        cm = {1: [np.zeros(0, dtype=np.int64)], 2: [np.zeros(0, dtype=np.int64)]}
        for fi, _, _, C_, _, _ in picked:
            if fi["role"]:
                cm[fi["role"]].append(C_.ravel())
        self.emap, self.imap = np.unique(np.concatenate(cm[1])), np.unique(np.concatenate(cm[2]))
        self.dims_full = (d["nvars"], d["neq"], d["niq"])
        nv, ne, ni = int(self.vmap.size), int(self.emap.size), int(self.imap.size)
        self.orc = Oracle(nv, ne, ni)
        rows = {1: [], 2: []}
        self.applications = 0
        for fi, desc, V, C_, ac, Cm in picked:
            Vc = np.searchsorted(self.vmap, V).astype(np.int32)
            Cc = None
            if fi["role"]:
                Cc = np.searchsorted(self.emap if fi["role"] == 1 else self.imap, C_).astype(np.int32)
            self.orc.add_function(fi["role"], desc, fi["linear"], Vc, Cc, ac)
            self.applications += int(V.shape[0])
            if fi["role"] and fi["unique_constraints"]:
                rows[fi["role"]].append(Cm.ravel())
        self.orc.analyze()
        vrows = np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in self.windows])
        if len(skip_vars):
            vrows = vrows[~np.isin(vrows, np.asarray(list(skip_vars), dtype=np.int64))]
        empty = np.zeros(0, dtype=np.int64)
        # the product's CSR rows of the window variables against the sub-problem's: same columns (after the renumbering)
        ip, ix = pattern if pattern is not None else nlp.pattern()
        sp, sx, _ = self.orc.pattern()
        PV, EC, IC = d["nvars"], d["neq"], d["niq"]
        SV = IC

        def compact_cols(cols):  # whole-problem KKT column -> sub-problem column (-1: not in the sub-problem)
            cols = np.asarray(cols, dtype=np.int64)
            out = np.full(cols.shape, -1, dtype=np.int64)
            for lo_c, hi_c, mp, base in ((0, PV, self.vmap, 0), (PV, PV + SV, self.imap, nv), (PV + SV, PV + SV + EC, self.emap, nv + ni),
                                         (PV + SV + EC, PV + SV + EC + IC, self.imap, nv + ni + ne)):
                sel = (cols >= lo_c) & (cols < hi_c)
                if sel.any():
                    loc = cols[sel] - lo_c
                    pos = np.searchsorted(mp, loc)
                    ok = (pos < mp.size) & (mp[np.minimum(pos, max(mp.size - 1, 0))] == loc) if mp.size else np.zeros(loc.shape, dtype=bool)
                    res = np.where(ok, pos + base, -1)
                    out[sel] = res
            return out
        kk, ks = [], []
        vr_c = np.searchsorted(self.vmap, vrows)
        for r, rc in zip(vrows, vr_c):
            a0, a1, b0, b1 = int(ip[r]), int(ip[r + 1]), int(sp[rc]), int(sp[rc + 1])
            if a1 - a0 != b1 - b0 or not np.array_equal(compact_cols(ix[a0:a1]), sx[b0:b1]):
                raise AssertionError(f"CSR row {r} of the product and of the oracle window differ in pattern")
            kk.append(np.arange(a0, a1, dtype=np.int64))
            ks.append(np.arange(b0, b1, dtype=np.int64))
        self.index = {"KKT": np.concatenate(kk) if kk else empty, "AGX": vrows, "PGX": vrows,
                      "FXE": np.concatenate(rows[1]) if rows[1] else empty, "FXI": np.concatenate(rows[2]) if rows[2] else empty}
        self._ref_index = {"KKT": np.concatenate(ks) if ks else empty, "AGX": vr_c, "PGX": vr_c,
                           "FXE": np.searchsorted(self.emap, self.index["FXE"]), "FXI": np.searchsorted(self.imap, self.index["FXI"])}

    def extract(self, out):
        """The looked-at elements of full-length host arrays."""
        return {k: np.asarray(out[k])[self.index[k]] for k in ARRAYS if k in out}

    def compare(self, slices, obj_scale, X, LE, LI, solver_coeffs=None, threads=0):
        """Worst relative difference per array (relative to the max-norm of the array's window elements)."""
        PV, EC, IC = self.dims_full
        SV = IC
        Xc = np.asarray(X, dtype=np.float64)[self.vmap]
        LEc = np.asarray(LE, dtype=np.float64)[self.emap] if self.emap.size else np.zeros(0)
        LIc = np.asarray(LI, dtype=np.float64)[self.imap] if self.imap.size else np.zeros(0)
        sc = None
        if solver_coeffs is not None:
            # NonLinearProgram.cpp:198-265 order: slack ones, primal diagonal, slack diagonal, equality pivots, inequality pivots
            s = np.asarray(solver_coeffs, dtype=np.float64)
            sc = np.concatenate([s[self.imap], s[SV + self.vmap], s[SV + PV + self.imap], s[SV + PV + SV + self.emap],
                                 s[SV + PV + SV + EC + self.imap]])
        ref = self.orc.eval("KKT", obj_scale, Xc, LEc, LIc, sc, threads=threads or hardware_threads())
        worst = {}
        for k, got in slices.items():
            want = ref[k][self._ref_index[k]]
            got = np.asarray(got)
            if want.size == 0:
                continue
            worst[k] = float(np.max(np.abs(got - want))) / max(float(np.max(np.abs(want))), 1e-300)
        worst["rows"] = int(self.index["AGX"].size)
        worst["applications"] = self.applications
        return worst


def check_windows(nlp, out, windows, obj_scale, X, LE, LI, solver_coeffs=None, pattern=None, threads=0):
    wc = WindowCheck(nlp, windows, pattern)
    return wc.compare(wc.extract(out), obj_scale, X, LE, LI, solver_coeffs, threads)
