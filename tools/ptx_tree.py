#!/usr/bin/env python
"""Canonical float expression trees of one PTX function (CPU-side parity aid).

Walks the function linearly, keeps reg -> expression, and prints a short hash (and optionally the
tree) of every value that is stored, returned or compared.  Leaves (loads, params, integer
conversions) are all 'L', so two functions with different argument layouts but the same arithmetic
give the same list.  Multiplicands of mul/fma and operands of add/min/max are sorted, so only a
genuinely different contraction (which product is fused, which is rounded) changes a hash.
usage: ptx_tree.py FILE.ptx FUNCTION_SUBSTRING [-v]
"""
import hashlib
import re
import sys

sys.path.insert(0, __file__.rsplit("/", 1)[0])
from ptx_ops import function_body  # noqa: E402

COMM = {"mul", "add", "max", "min"}


def h(s):
    return hashlib.md5(s.encode()).hexdigest()[:8]


def trees(body, verbose=False):
    env = {}
    out = []

    def val(tok):
        tok = tok.strip()
        if tok in env:
            return env[tok]
        if re.match(r"0[fF][0-9A-Fa-f]{8}$", tok):
            return "C" + tok[2:].upper()
        return "L"

    ndef = {}
    for line in body.splitlines():
        m = re.match(r"\s*(@!?%p\d+\s+)?([a-z0-9_.]+)\s+(%f\d+),", line)
        if m:
            ndef[m.group(3)] = ndef.get(m.group(3), 0) + 1
    for line in body.splitlines():
        line = line.strip()
        m = re.match(r"(@!?%p\d+\s+)?([a-z0-9_.]+)\s+(.*);", line)
        if not m:
            continue
        op, args = m.group(2), [a.strip() for a in m.group(3).split(",")]
        base = op.split(".")[0]
        if ".f32" not in op and base not in ("mov", "selp"):
            continue
        if args and ndef.get(args[0], 0) > 1 and base not in ("st", "setp"):
            # a register defined on several paths (a phi): report each incoming value
            pending_phi = args[0]
        else:
            pending_phi = None
        if base in ("st",):
            for tok in re.findall(r"%f\d+|0[fF][0-9A-Fa-f]{8}", m.group(3).split("]", 1)[1]):
                out.append(("store", val(tok)))
        elif base == "setp":
            cmpop = op.split(".")[1]
            out.append(("cmp." + cmpop, "(" + val(args[1]) + "," + val(args[2]) + ")"))
        elif base in ("ld", "cvt"):
            for tok in re.findall(r"%f\d+", m.group(3).split("[", 1)[0]):
                env[tok] = "L"
        elif base == "mov":
            if args[0].startswith("%f"):
                env[args[0]] = val(args[1])
        elif base == "selp":
            if args[0].startswith("%f"):
                env[args[0]] = "sel(" + val(args[1]) + "," + val(args[2]) + ")"
        elif base == "fma":
            a, b = sorted([val(args[1]), val(args[2])])
            env[args[0]] = f"fma({a},{b},{val(args[3])})"
        elif base == "call":
            pass
        else:
            vs = [val(a) for a in args[1:]]
            if base in COMM:
                vs.sort()
            env[args[0]] = base + "(" + ",".join(vs) + ")"
        if pending_phi and pending_phi in env:
            out.append(("phi", env[pending_phi]))
    return out


if __name__ == "__main__":
    text = open(sys.argv[1]).read()
    verbose = "-v" in sys.argv
    for kind, e in trees(function_body(text, sys.argv[2])):
        print(kind, h(e), len(e), e[:400] if verbose else "")
