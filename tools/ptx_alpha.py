#!/usr/bin/env python
"""Alpha-rename the float registers of one PTX function in order of first appearance and print its
float instructions, so that two builds of the same arithmetic can be compared with `diff`.
usage: ptx_alpha.py FILE.ptx FUNCTION_SUBSTRING"""
import re
import sys

sys.path.insert(0, __file__.rsplit("/", 1)[0])
from ptx_ops import function_body  # noqa: E402

body = function_body(open(sys.argv[1]).read(), sys.argv[2])
names = {}


def ren(m):
    return names.setdefault(m.group(0), "f%d" % len(names))


for line in body.splitlines():
    line = line.strip()
    if re.match(r"(@\S+\s+)?(mul|add|sub|fma|div|sqrt|rcp|max|min|neg|abs|setp|selp|mov)\.[a-z0-9.]*f32", line):
        line = re.sub(r"%f\d+", ren, line)
        line = re.sub(r"%p\d+", "p", line).replace("mul.rn.f32", "mul.f32")
        line = re.sub(r"\s+", " ", line).rstrip(";")
        op, rest = line.split(" ", 1) if not line.startswith("@") else (" ".join(line.split(" ")[:2]), line.split(" ", 2)[2])
        a = [x.strip() for x in rest.split(",")]
        base = op.split()[-1].split(".")[0]
        if base in ("mul", "add", "max", "min") and len(a) == 3:
            a[1:] = sorted(a[1:])
        if base == "fma":
            a[1:3] = sorted(a[1:3])
        print(op, ", ".join(a))
