#!/usr/bin/env python
"""Checks that the assembler kept the program order of the parked-value accesses of a specialised kernel.

The specialiser parks values in lane-private shared-memory slots and recycles the slots (csrc/jit.cpp).  ptxas 12.9
was caught sinking the store of a recycled slot below the later load of the same address (LGL7 fused piece with plain
st.shared / ld.shared), which makes the reload return the slot's previous value.  For every slot address this script
compares the store / load sequence of the PTX with the one in the SASS of the assembled cubin: loads between two stores
may be permuted, nothing may cross a store.

    python tools/sass_park_order.py kernel.ptx        # assembles with ptxas, disassembles with cuobjdump
"""
import collections
import os
import re
import subprocess
import sys
import tempfile


def ptx_sequence(ptx):
    seq = collections.defaultdict(list)
    for line in ptx.splitlines():
        m = re.match(r"st(?:\.volatile)?\.shared\.f64 \[%rsb(?:\+(\d+))?\]", line)
        if m:
            seq[int(m.group(1) or 0)].append("S")
            continue
        m = re.match(r"ld(?:\.volatile)?\.shared\.f64 %fd\d+, \[%rsb(?:\+(\d+))?\]", line)
        if m:
            seq[int(m.group(1) or 0)].append("L")
    return seq


def sass_sequence(sass):
    seq = collections.defaultdict(list)
    for line in sass.splitlines():
        m = re.search(r"\b(STS|LDS)(?:\.[A-Z0-9]+)*\.64 (?:R\d+, )?\[R\d+(?:\+0x([0-9a-f]+))?\]", line)
        if m:
            seq[int(m.group(2) or "0", 16)].append("S" if m.group(1) == "STS" else "L")
    return seq


def check(ptx_text):
    """Returns a list of (address, ptx order, sass order) for every slot whose accesses were reordered across a store."""
    with tempfile.TemporaryDirectory() as d:
        p, c = os.path.join(d, "k.ptx"), os.path.join(d, "k.cubin")
        open(p, "w").write(ptx_text)
        subprocess.check_call(["ptxas", "-arch=sm_100a", "-O3", p, "-o", c])
        sass = subprocess.check_output(["cuobjdump", "-sass", c], text=True)
    a, b = ptx_sequence(ptx_text), sass_sequence(sass)
    bad = []
    for addr, s in a.items():
        if "".join(s) != "".join(b.get(addr, [])):
            bad.append((addr, "".join(s), "".join(b.get(addr, []))))
    return bad


if __name__ == "__main__":
    rc = 0
    for f in sys.argv[1:]:
        bad = check(open(f).read())
        print(f, "reordered slots:", bad if bad else "none")
        rc |= bool(bad)
    sys.exit(rc)
