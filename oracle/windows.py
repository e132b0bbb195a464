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
        problem keeps applications of its neighbours too; their rows live on the neighbour)."""
        d = nlp.dims()
        self.windows = [(int(a), int(b)) for a, b in windows]
        self.orc = Oracle(d["nvars"], d["neq"], d["niq"])
        lo = np.array([w[0] for w in self.windows], dtype=np.int64)
        hi = np.array([w[1] for w in self.windows], dtype=np.int64)
        rows = {1: [], 2: []}
        self.applications = 0
        for f in range(d["nfuncs"]):
            fi = nlp.function_info(f)
            V, Cc = nlp.function_indices(f)
            hit = np.zeros(V.shape[0], dtype=bool)
            for a, b in zip(lo, hi):
                hit |= ((V >= a) & (V < b)).any(axis=1)
            if not hit.any():
                continue
            ac = nlp.function_appconst(f)
            self.orc.add_function(fi["role"], nlp.tape_describe(fi["tape"]), fi["linear"], V[hit], Cc[hit] if fi["role"] else None,
                                  ac[hit] if ac.size else None)
            self.applications += int(hit.sum())
            if fi["role"] and fi["unique_constraints"]:
                mine = hit.copy()
                if own_apps is not None:
                    lo_a, hi_a = own_apps(f)
                    mine[:lo_a] = False
                    mine[hi_a:] = False
                rows[fi["role"]].append(Cc[mine].ravel().astype(np.int64))
        self.orc.analyze()
        vrows = np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in self.windows])
        if len(skip_vars):
            vrows = vrows[~np.isin(vrows, np.asarray(list(skip_vars), dtype=np.int64))]
        empty = np.zeros(0, dtype=np.int64)
        # the product's CSR rows of the window variables against the sub-problem's: same columns, bit for bit
        ip, ix = pattern if pattern is not None else nlp.pattern()
        sp, sx, _ = self.orc.pattern()
        kk, ks = [], []
        for r in vrows:
            a0, a1, b0, b1 = int(ip[r]), int(ip[r + 1]), int(sp[r]), int(sp[r + 1])
            if a1 - a0 != b1 - b0 or not np.array_equal(ix[a0:a1], sx[b0:b1]):
                raise AssertionError(f"CSR row {r} of the product and of the oracle window differ in pattern")
            kk.append(np.arange(a0, a1, dtype=np.int64))
            ks.append(np.arange(b0, b1, dtype=np.int64))
        self.index = {"KKT": np.concatenate(kk) if kk else empty, "AGX": vrows, "PGX": vrows,
                      "FXE": np.concatenate(rows[1]) if rows[1] else empty, "FXI": np.concatenate(rows[2]) if rows[2] else empty}
        self._ref_index = dict(self.index)
        self._ref_index["KKT"] = np.concatenate(ks) if ks else empty

    def extract(self, out):
        """The looked-at elements of full-length host arrays."""
        return {k: np.asarray(out[k])[self.index[k]] for k in ARRAYS if k in out}

    def compare(self, slices, obj_scale, X, LE, LI, solver_coeffs=None, threads=0):
        """Worst relative difference per array (relative to the max-norm of the array's window elements)."""
        ref = self.orc.eval("KKT", obj_scale, X, LE, LI, solver_coeffs, threads=threads or hardware_threads())
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
