"""Test-only numpy interpreter of the device tape format (asset_asrl_b200/csrc/tape.h).

It lets the CPU test-suite check the host lowering (expression -> scalar DAG -> symbolic adjoint -> tape with slot
allocation) without a GPU: the same forward-mode dual semantics as kernels.cu, all tangent directions at once.
Never imported by the product.
"""
import numpy as np

from asset_asrl_b200.VectorFunctions import OP


def run_tape(info, code, x, lam, appconst=(), full=True):
    """Returns f [nout], J [nout, nin], g [nin], H [nin, nin] (g, H None when the adjoint section is not run)."""
    nin, nout = info["nin"], info["nout"]
    nins = info["n_instr"] if (full and info["has_adjoint"]) else info["n_primal"]
    ns = max(info["slots_total"], 1)
    v = np.zeros(ns)
    d = np.zeros((ns, nin))
    f = np.zeros(nout)
    J = np.zeros((nout, nin))
    g = np.zeros(nin)
    H = np.zeros((nin, nin))
    cbits = code["imm"].view("<u8")
    x = np.asarray(x, dtype=np.float64)
    for pc in range(nins):
        op, dst, a, b = int(code["op"][pc]), int(code["dst"][pc]), int(code["a"][pc]), int(code["b"][pc])
        imm = float(code["imm"][pc])
        name = [k for k, val in OP.items() if val == op][0]
        if name == "OUT_F":
            f[dst] = v[a]
            J[dst] = d[a]
            continue
        if name == "OUT_G":
            g[dst] = v[a]
            H[dst] = d[a]
            continue
        va, da = (v[a], d[a]) if a < ns else (0.0, 0.0)
        vb, db = (v[b], d[b]) if b < ns else (0.0, 0.0)
        with np.errstate(all="ignore"):
            if name == "CONST":
                rv, rd = imm, np.zeros(nin)
            elif name == "INPUT":
                rv, rd = x[a], np.eye(nin)[a]
            elif name == "LAMBDA":
                rv, rd = lam[a], np.zeros(nin)
            elif name == "APPCONST":
                rv, rd = appconst[a], np.zeros(nin)
            elif name == "ADD":
                rv, rd = va + vb, da + db
            elif name == "SUB":
                rv, rd = va - vb, da - db
            elif name == "MUL":
                rv, rd = va * vb, da * vb + va * db
            elif name == "DIV":
                rv = va / vb
                rd = (da - rv * db) / vb
            elif name == "NEG":
                rv, rd = -va, -da
            elif name == "ADDC":
                rv, rd = va + imm, da
            elif name == "MULC":
                rv, rd = va * imm, da * imm
            elif name == "SQUARE":
                rv, rd = va * va, 2 * va * da
            elif name == "INV":
                rv = 1 / va
                rd = -da * rv * rv
            elif name == "SQRT":
                rv = np.sqrt(va)
                rd = 0.5 * da / rv
            elif name == "EXP":
                rv = np.exp(va)
                rd = rv * da
            elif name == "LOG":
                rv, rd = np.log(va), da / va
            elif name == "SIN":
                rv, rd = np.sin(va), np.cos(va) * da
            elif name == "COS":
                rv, rd = np.cos(va), -np.sin(va) * da
            elif name == "TAN":
                rv = np.tan(va)
                rd = da * (1 + rv * rv)
            elif name == "ASIN":
                rv, rd = np.arcsin(va), da / np.sqrt(1 - va * va)
            elif name == "ACOS":
                rv, rd = np.arccos(va), -da / np.sqrt(1 - va * va)
            elif name == "ATAN":
                rv, rd = np.arctan(va), da / (1 + va * va)
            elif name == "SINH":
                rv, rd = np.sinh(va), np.cosh(va) * da
            elif name == "COSH":
                rv, rd = np.cosh(va), np.sinh(va) * da
            elif name == "TANH":
                rv = np.tanh(va)
                rd = da * (1 - rv * rv)
            elif name == "ASINH":
                rv, rd = np.arcsinh(va), da / np.sqrt(va * va + 1)
            elif name == "ACOSH":
                rv, rd = np.arccosh(va), da / np.sqrt(va * va - 1)
            elif name == "ATANH":
                rv, rd = np.arctanh(va), da / (1 - va * va)
            elif name == "ABS":
                rv, rd = abs(va), np.sign(va) * da
            elif name == "SIGN":
                rv, rd = np.sign(va), np.zeros(nin)
            elif name == "POWC":
                rv, rd = va ** imm, imm * va ** (imm - 1) * da
            elif name == "ATAN2":
                rv = np.arctan2(va, vb)
                rd = (vb * da - va * db) / (va * va + vb * vb)
            elif name == "LT":
                rv, rd = float(va < vb), np.zeros(nin)
            elif name == "LE":
                rv, rd = float(va <= vb), np.zeros(nin)
            elif name == "SELECT":
                c = int(cbits[pc])
                rv, rd = (vb, db) if va != 0 else (v[c], d[c])
            elif name == "FMA":
                c = int(cbits[pc])
                rv, rd = va * vb + v[c], da * vb + va * db + d[c]
            else:
                raise ValueError(f"unknown op {op}")
        v[dst] = rv
        d[dst] = rd
    if nins == info["n_primal"]:
        return f, J, None, None
    return f, J, g, H
