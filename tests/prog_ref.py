"""Test-only numpy interpreter of the device program format (asset_asrl_b200/csrc/expr.hpp: Program / PInstr).

Same semantics as k_eval_prog in kernels.cu for ONE application, so the CPU test-suite can check the whole host
lowering (expression -> scalar DAG -> symbolic adjoint -> sparse second-order sweep -> FMA fusion -> bundling -> slot
allocation) without a GPU.  Never imported by the product.
"""
import math

import numpy as np

from asset_asrl_b200.VectorFunctions import OP

PK_FMA, PK_LOAD, PK_OUT, PK_SPECIAL, PK_END = range(5)
PO_F, PO_J, PO_G, PO_H = range(4)
_NAME = {v: k for k, v in OP.items()}


def _interval(axis, t):
    """OP_TABIDX (csrc/tape.h): axis = [n, evenly spaced, coordinates...]."""
    n, ts = int(axis[0]), axis[2:2 + int(axis[0])]
    if axis[1] != 0.0:
        e = int(np.trunc(np.clip((t - ts[0]) / (ts[1] - ts[0]), -1.0, float(n))))
    else:
        e = int(np.searchsorted(ts, t, side="right")) - 1
    return float(max(min(e, n - 2), 0))


def _special(op, a, b, c, imm, tables=None):
    n = _NAME[op]
    if n == "TABIDX":
        return _interval(tables[int(imm):], a)
    if n == "TABLOAD":
        return float(tables[int(imm) + int(a)])
    with np.errstate(all="ignore"):
        return {
            "DIV": lambda: a / b, "INV": lambda: 1.0 / a, "SQRT": lambda: math.sqrt(a), "EXP": lambda: math.exp(a),
            "LOG": lambda: math.log(a), "SIN": lambda: math.sin(a), "COS": lambda: math.cos(a), "TAN": lambda: math.tan(a),
            "ASIN": lambda: math.asin(a), "ACOS": lambda: math.acos(a), "ATAN": lambda: math.atan(a),
            "SINH": lambda: math.sinh(a), "COSH": lambda: math.cosh(a), "TANH": lambda: math.tanh(a),
            "ASINH": lambda: math.asinh(a), "ACOSH": lambda: math.acosh(a), "ATANH": lambda: math.atanh(a),
            "ABS": lambda: abs(a), "SIGN": lambda: float(np.sign(a)), "POWC": lambda: a ** imm,
            "ATAN2": lambda: math.atan2(a, b), "LT": lambda: float(a < b), "LE": lambda: float(a <= b),
            "SELECT": lambda: b if a != 0.0 else c,
        }[n]()


def run_program(stats, code, consts, x, lam, appconst=(), tables=None):
    """Returns (F dict k->v, G dict i->v, T dict triplet_offset->v) of one application."""
    slots = np.zeros(stats["slots"] + 1)

    def val(w):
        w = int(w)
        v = slots[w >> 3] if (w & 1) else consts[w >> 3]
        return -v if (w & 2) else v

    F, G, T = {}, {}, {}
    nb = len(code) // 4
    for b in range(nb):
        ins = code[4 * b:4 * b + 4]
        head = int(ins[0]["d"])
        kind, cnt = head >> 28, (head >> 24) & 15
        if kind == PK_END:
            break
        if kind == PK_FMA:
            vals = [val(ins[q]["a"]) * val(int(ins[q]["b"]) & ~2) + val(ins[q]["c"]) for q in range(cnt)]
            for q in range(cnt):
                slots[(int(ins[q]["d"]) & 0x00FFFFFF) >> 3] = vals[q]
        elif kind == PK_LOAD:
            for q in range(cnt):
                idx, ty = int(ins[q]["a"]), int(ins[q]["b"])
                v = x[idx] if ty == 0 else (lam[idx] if ty == 1 else appconst[idx])
                slots[(int(ins[q]["d"]) & 0x00FFFFFF) >> 3] = v
        elif kind == PK_OUT:
            for q in range(cnt):
                v, ok, idx = val(ins[q]["a"]), int(ins[q]["b"]), int(ins[q]["c"])
                (F if ok == PO_F else G if ok == PO_G else T)[idx] = v
        elif kind == PK_SPECIAL:
            w = int(ins[0]["c"])
            r = _special(w & 0xFF, val(ins[0]["a"]), val(ins[0]["b"]), val(ins[1]["a"]), consts[w >> 8], tables)
            slots[(int(ins[0]["d"]) & 0x00FFFFFF) >> 3] = r
        else:
            raise ValueError(f"bad bundle kind {kind}")
    return F, G, T


def dense_blocks(fi, F, G, T):
    """fx [or], jx [or][ir], gx [ir] (zeros when not produced), hx [ir][ir] from the program outputs of one application."""
    ir, orr = fi["ir"], fi["orows"]
    lay_h, lay_j = not fi["linear"], fi["role"] != 0
    fx = np.array([F.get(k, 0.0) for k in range(orr)])
    gx = np.array([G.get(i, 0.0) for i in range(ir)])
    jx, hx = np.zeros((orr, ir)), np.zeros((ir, ir))
    e = 0
    for i in range(ir):
        if lay_h:
            for j in range(i, ir):
                hx[j, i] = hx[i, j] = T.get(e, 0.0)
                e += 1
        if lay_j:
            for k in range(orr):
                jx[k, i] = T.get(e, 0.0)
                e += 1
    return fx, jx, gx, hx
