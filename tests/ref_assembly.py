"""Test-only CPU assembly of the KKT outputs from per-application dense blocks.

Given a per-application evaluator (the numpy tape interpreter or the C oracle) it performs the reference's
scatter ``values[KKTLocations[k]] += v`` in canonical order (DenseFunctionBase.h:1413-1466) so that GPU results can be
compared entry by entry.  Small problems only.
"""
import numpy as np

from prog_ref import run_program, dense_blocks


def tape_block_eval(nlp):
    """Per-application evaluator running the unlinked device program (mode 4 = values + Jacobian + gradient + Hessian)."""
    progs = {}

    def ev(f, fi, x, lam, appc):
        if f not in progs:
            progs[f] = nlp.program(f, 4) + (nlp.program_tables(f, 4),)
        st, code, consts, tables = progs[f]
        F, G, T = run_program(st, code, consts, x, lam, appc, tables=tables)
        fx, jx, gx, hx = dense_blocks(fi, F, G, T)
        if fi["role"] == 0:  # objectives scatter no Jacobian triplets: recover the gradient row from G (= lam * df/dx)
            jx = (gx / lam[0]).reshape(1, -1) if lam[0] != 0 else jx
        return fx, jx, hx

    return ev


def assemble(nlp, X, LE, LI, obj_scale=1.0, block_eval=None, solver_coeffs=None):
    """Returns dict(val, PGX, AGX, FXE, FXI, KKT) following NonLinearProgram::evalKKT."""
    d = nlp.dims()
    ev = block_eval or tape_block_eval(nlp)
    loc = nlp.kkt_locations()
    KKT = np.zeros(d["nnz"])
    PGX, AGX = np.zeros(d["nvars"]), np.zeros(d["nvars"])
    FXE, FXI = np.zeros(d["neq"]), np.zeros(d["niq"])
    val = 0.0
    k = 0
    order = [f for role in (0, 1, 2) for f in range(d["nfuncs"]) if nlp.function_info(f)["role"] == role]
    for f in order:
        fi = nlp.function_info(f)
        V, Cc = nlp.function_indices(f)
        L = LE if fi["role"] == 1 else LI
        appc_all = nlp.function_appconst(f)
        for a in range(fi["nappl"]):
            x = X[V[a]]
            lam = np.array([obj_scale]) if fi["role"] == 0 else L[Cc[a]]
            appc = appc_all[a] if fi["nappconst"] else ()
            fx, J, H = ev(f, fi, x, lam, appc)
            if fi["role"] == 0:
                val += fx[0] * obj_scale
                np.add.at(PGX, V[a], J[0] * obj_scale)
            else:
                (FXE if fi["role"] == 1 else FXI)[Cc[a]] += fx
            g = J.T @ lam
            if fi["role"] != 0:
                np.add.at(AGX, V[a], g)
            for i in range(fi["ir"]):
                if not fi["linear"]:
                    for j in range(i, fi["ir"]):
                        KKT[loc[k]] += H[j, i]
                        k += 1
                if fi["role"] != 0:
                    for j in range(fi["orows"]):
                        KKT[loc[k]] += J[j, i]
                        k += 1
    assert k == d["n_user"], (k, d["n_user"])
    if solver_coeffs is not None:
        np.add.at(KKT, loc[d["n_user"]:], solver_coeffs)
    return dict(val=val, PGX=PGX, AGX=AGX, FXE=FXE, FXI=FXI, KKT=KKT)
