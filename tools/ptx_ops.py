#!/usr/bin/env python
"""Print the floating-point dataflow of one PTX function as canonical expression trees.

Used on the CPU to check that a BSDF entry point of this library contracts (fma vs mul+add) the
same way as the reference's function of the same role: compile both to PTX and diff the output.
usage: ptx_ops.py FILE.ptx FUNCTION_SUBSTRING
"""
import re
import sys


def function_body(text, name):
    # definitions (with a body) come after the forward declarations
    starts = [m.start() for m in re.finditer(r"^\.(?:visible \.)?(?:func|entry)[^\n]*" + re.escape(name), text, re.M)]
    for s in reversed(starts):
        brace = text.find("{", s)
        semi = text.find(";", s)
        if brace != -1 and (semi == -1 or brace < semi or text[s:brace].count("(") > 0 and text[s:brace].rstrip().endswith(")")):
            end = text.find("\n}", brace)
            return text[brace:end]
    raise SystemExit("function not found: " + name)


FOPS = ("mul", "add", "sub", "fma", "div", "sqrt", "rcp", "max", "min", "neg", "abs", "rsqrt", "cvt", "sin", "cos", "ex2", "lg2")


def ops(body):
    out = []
    for line in body.splitlines():
        line = line.strip()
        m = re.match(r"(@!?%p\d+\s+)?([a-z0-9_.]+)\s+(.*);", line)
        if not m:
            continue
        op = m.group(2)
        if ".f32" in op and op.split(".")[0] in FOPS:
            out.append(op)
    return out


if __name__ == "__main__":
    text = open(sys.argv[1]).read()
    for o in ops(function_body(text, sys.argv[2])):
        print(o)
