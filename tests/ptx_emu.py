"""CPU emulation of the straight-line PTX the kernel specialiser emits (csrc/jit.cpp, affine variant with direct stores).

Test infrastructure: lets the CPU suite execute a specialised / fused kernel for single threads (one `rel` each) without
a GPU and without the assembler, so that a wrong result on the device can be attributed to the emitter (reproduced
here) or to the assembler (not reproduced).  Only the instruction subset jit.cpp generates is understood.
"""
import math
import re
import struct

import numpy as np

PARAMS = ["pX", "pL", "pK", "pF", "pS", "pA", "pVI", "pCI", "pKD", "pFD"]

_SPECIAL = {
    "ab_sp_exp": math.exp, "ab_sp_log": math.log, "ab_sp_sin": math.sin, "ab_sp_cos": math.cos, "ab_sp_tan": math.tan,
    "ab_sp_asin": math.asin, "ab_sp_acos": math.acos, "ab_sp_atan": math.atan, "ab_sp_sinh": math.sinh,
    "ab_sp_cosh": math.cosh, "ab_sp_tanh": math.tanh, "ab_sp_asinh": math.asinh, "ab_sp_acosh": math.acosh,
    "ab_sp_atanh": math.atanh, "ab_sp_pow": math.pow, "ab_sp_atan2": math.atan2,
}


def _imm(tok):
    return struct.unpack(">d", bytes.fromhex(tok[2:]))[0]


class Kernel:
    """Parsed kernel: address groups and the instruction list of the body."""

    def __init__(self, ptx):
        # interpolation tables: module-scope data the table ops address (jit.cpp: finish())
        m = re.search(r"\.global \.align 16 \.b64 ab_tab\[(\d+)\] = \{([^}]*)\}", ptx)
        self.table = None
        if m:
            words = [int(w, 16) for w in m.group(2).replace("\n", " ").split(",")]
            assert len(words) == int(m.group(1))
            self.table = np.array(words, dtype="<u8").view("<f8")
        body = ptx[ptx.index(".visible .entry"):]
        self.lines = [ln.strip() for ln in body.split("\n")]
        self.groups = {}  # %rgN -> (param index, byte stride per rel, byte offset)
        self.sector = {}  # %rhN -> (param index, byte stride per rel, byte offset, multiplier of the alignment term)
        pend = pend_m = None
        for ln in self.lines:
            m = re.match(r"mul\.wide\.s32 %rt, %rel, (-?\d+);", ln)
            if m:
                pend = int(m.group(1))
                continue
            m = re.match(r"add\.s64 %rg(\d+), %rt, %rp(\d+);", ln)
            if m:
                self.groups[int(m.group(1))] = [int(m.group(2)), pend, 0]
                continue
            m = re.match(r"add\.s64 %rg(\d+), %rg\1, (-?\d+);", ln)
            if m:
                self.groups[int(m.group(1))][2] += int(m.group(2))
                continue
            # sector base registers: %rh<g> = array + rel * stride + constant + ((-(rel * m)) mod 4) * 8
            m = re.match(r"add\.s64 %rh(\d+), %rt, %rp(\d+);", ln)
            if m:
                self.sector[int(m.group(1))] = [int(m.group(2)), pend, 0, 0]
                continue
            m = re.match(r"add\.s64 %rh(\d+), %rh\1, (-?\d+);", ln)
            if m:
                self.sector[int(m.group(1))][2] += int(m.group(2))
                continue
            m = re.match(r"mul\.lo\.s32 %ri, %rel, (\d+);", ln)
            if m:
                pend_m = int(m.group(1))
                continue
            m = re.match(r"add\.s64 %rh(\d+), %rh\1, %rt;", ln)
            if m:
                self.sector[int(m.group(1))][3] = pend_m

    def run(self, arrays, rel, scale=1.0):
        """Executes thread `rel`; `arrays` maps parameter index (see PARAMS) -> float64 numpy array (modified in place)."""
        R = {"%fscale": float(scale), "%fzero": 0.0}
        shared = {}
        preds = {"%pm2_0": rel % 2 == 0, "%pm2_1": rel % 2 == 1}
        for r in range(4):
            preds["%pm4_" + str(r)] = rel % 4 == r

        def val(tok):
            tok = tok.strip().rstrip(";").rstrip(",")
            if tok.startswith("0d"):
                return _imm(tok)
            return R[tok]

        def gaddr(expr):
            m = re.match(r"\[%r([gh])(\d+)(?:\+(-?\d+))?\]", expr)
            if m.group(1) == "g":
                par, st, off = self.groups[int(m.group(2))]
                b = st * rel + off + int(m.group(3) or 0)
            else:
                par, st, off, mul = self.sector[int(m.group(2))]
                b = st * rel + off + ((-(rel * mul)) & 3) * 8 + int(m.group(3) or 0)
            assert b % 8 == 0
            return par, b // 8, b

        def saddr(expr):
            m = re.match(r"\[%rsb(?:\+(\d+))?\]", expr)
            return int(m.group(1) or 0)

        for ln in self.lines:
            if not ln or ln in ("{", "}") or ln.startswith((".reg", "//", ".visible", ".maxntid", "L_done", "ret")):
                continue
            pred = True
            if ln.startswith("@"):
                p, ln = ln.split(" ", 1)
                neg = p.startswith("@!")
                name = p[2:] if neg else p[1:]
                if name not in preds:
                    continue  # prologue predicates (%p0: thread past the end)
                pred = preds[name] != neg
            op, _, rest = ln.partition(" ")
            rest = rest.rstrip(";")
            if op == "fma.rn.f64":
                d, a, b, c = [t.strip() for t in rest.split(",")]
                R[d] = float(np.float64(math.fma(val(a), val(b), val(c)))) if hasattr(math, "fma") else _fma(val(a), val(b), val(c))
            elif op in ("neg.f64", "abs.f64", "mov.f64", "rcp.rn.f64", "sqrt.rn.f64"):
                d, a = [t.strip() for t in rest.split(",")]
                x = val(a)
                R[d] = {"neg.f64": -x, "abs.f64": abs(x), "mov.f64": x, "rcp.rn.f64": (1.0 / x) if x != 0 else math.inf,
                        "sqrt.rn.f64": math.sqrt(x) if x >= 0 else math.nan}[op]
            elif op in ("div.rn.f64", "mul.rn.f64"):
                d, a, b = [t.strip() for t in rest.split(",")]
                R[d] = val(a) / val(b) if op == "div.rn.f64" else val(a) * val(b)
            elif op.startswith("ld.global.nc"):
                m = re.match(r"(?:\{([^}]*)\}|(%fd\d+)), (\[.*\])", rest)
                regs = [t.strip() for t in (m.group(1) or m.group(2)).split(",")]
                par, el, byte = gaddr(m.group(3))
                assert byte % (8 * len(regs)) == 0, "misaligned vector load"
                for j, r in enumerate(regs):
                    R[r] = float(arrays[par][el + j])
            elif op.startswith("st.global"):
                m = re.match(r"(\[[^\]]*\]), (?:\{([^}]*)\}|(\S+))", rest)
                regs = [t.strip() for t in (m.group(2) or m.group(3)).split(",")]
                par, el, byte = gaddr(m.group(1))
                if pred:
                    assert byte % (8 * len(regs)) == 0, "misaligned vector store"
                    for j, r in enumerate(regs):
                        arrays[par][el + j] = val(r)
            elif op in ("st.shared.f64", "st.volatile.shared.f64"):
                a, v = [t.strip() for t in rest.split(",")]
                shared[saddr(a)] = val(v)
            elif op in ("ld.shared.f64", "ld.volatile.shared.f64"):
                d, a = [t.strip() for t in rest.split(",", 1)]
                R[d] = shared[saddr(a.strip())]
            elif op == "{" and ".reg" in rest:  # a table op: a self-contained block with its own registers
                self._block(rest, R, val)
            elif op == "{":  # call of a transcendental helper
                fn = re.search(r"call\.uni \(po\), (\w+)", rest).group(1)
                args = re.findall(r"st\.param\.f64 \[pi\d\], ([^;]+);", rest)
                d = re.search(r"ld\.param\.f64 (%fd\d+), \[po\]", rest).group(1)
                R[d] = _SPECIAL[fn](*[val(a) for a in args])
            elif op.startswith(("setp", "selp")):
                _cmp(op, rest, R, preds, val)
            elif op.startswith(("ld.param", "cvta", "mov.u32", "mad.", "add.s", "add.u", "and.b32", "mul.wide", "sub.s32", "min.s32",
                                "shl.", "cvt.", "bra", "bar.", "not.", "mov.b32", "mul.lo", "neg.s32")):
                continue  # prologue / address set-up, handled through the groups table
            else:
                raise NotImplementedError("ptx_emu: " + ln)


def _block(self, text, R, val):
    """OP_TABIDX / OP_TABLOAD as jit.cpp emits them: integer index arithmetic and loads from the module's table."""
    L, P = {}, {}

    def get(tok):
        tok = tok.strip()
        if tok in L:
            return L[tok]
        if re.fullmatch(r"-?\d+", tok):
            return int(tok)
        return val(tok)

    def put(tok, v):
        if tok in L:
            L[tok] = v
        else:
            R[tok] = v

    for st in text.strip().rstrip("}").split(";"):
        st = st.strip()
        if not st:
            continue
        op, _, rest = st.partition(" ")
        a = [t.strip() for t in rest.split(",")]
        if op == ".reg":
            L[rest.split()[-1]] = 0
        elif op == "mov.u64":
            assert a[1] == "ab_tab"
            L[a[0]] = 0  # byte offset into the table
        elif op == "mov.s32":
            L[a[0]] = get(a[1])
        elif op == "cvt.rzi.s32.f64":
            x = get(a[1])
            L[a[0]] = 0 if x != x else int(max(min(x, 2147483647.0), -2147483648.0))
        elif op == "mad.wide.s32":
            L[a[0]] = get(a[1]) * get(a[2]) + get(a[3])
        elif op == "ld.global.nc.f64":
            m = re.fullmatch(r"\[(%\w+)\+(-?\d+)\]", a[1])
            addr = L[m.group(1)] + int(m.group(2))
            assert addr % 8 == 0 and 0 <= addr // 8 < self.table.size, "table access out of range"
            put(a[0], float(self.table[addr // 8]))
        elif op in ("add.s32", "sub.s32", "min.s32", "max.s32"):
            x, y = get(a[1]), get(a[2])
            L[a[0]] = {"add.s32": x + y, "sub.s32": x - y, "min.s32": min(x, y), "max.s32": max(x, y)}[op]
        elif op == "setp.le.s32":
            P[a[0]] = get(a[1]) <= get(a[2])
        elif op == "setp.le.and.f64":
            P[a[0]] = (get(a[1]) <= get(a[2])) and P[a[3]]
        elif op == "selp.s32":
            L[a[0]] = get(a[1]) if P[a[3]] else get(a[2])
        elif op == "sub.rn.f64":
            put(a[0], get(a[1]) - get(a[2]))
        elif op == "div.rn.f64":
            put(a[0], get(a[1]) / get(a[2]))
        elif op == "cvt.rn.f64.s32":
            put(a[0], float(get(a[1])))
        else:
            raise NotImplementedError("ptx_emu block: " + st)


Kernel._block = _block


def _cmp(op, rest, R, preds, val):
    parts = [t.strip() for t in rest.split(",")]
    if op.startswith("setp") and op.endswith(".f64"):
        c = op.split(".")[1]
        a, b = val(parts[1]), val(parts[2])
        preds[parts[0]] = {"gt": a > b, "lt": a < b, "le": a <= b, "ge": a >= b, "neu": a != b or a != a or b != b, "eq": a == b}[c]
    elif op == "selp.f64":
        R[parts[0]] = val(parts[1]) if preds[parts[3]] else val(parts[2])


def _fma(a, b, c):
    from fractions import Fraction
    return float(Fraction(a) * Fraction(b) + Fraction(c))
