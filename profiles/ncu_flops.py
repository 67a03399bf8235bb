#!/usr/bin/env python
"""Executed FP32 flops of one kernel launch from an .ncu-rep source page: sum over SASS instructions of predicated-on thread
instructions x flops per instruction (FADD / FMUL = 1, FFMA = 2; the packed FADD2 / FMUL2 / FFMA2 of sm_100 count twice).
usage: python profiles/ncu_flops.py rep.ncu-rep [launch-index]"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], []
for row in csv.reader(out.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = []; blocks.append((row[1], cur))
    elif cur is not None and blocks:
        cur.append(row)
name, rows = blocks[which]
hdr = rows[0]
isrc, ipred = hdr.index("Source"), hdr.index("Predicated-On Thread Instructions Executed")
W = {"FADD": 1, "FMUL": 1, "FFMA": 2, "FADD2": 2, "FMUL2": 2, "FFMA2": 4}
flops, by = 0, {}
for r in rows[1:]:
    try:
        src, n = r[isrc].split(), int(r[ipred])
    except (ValueError, IndexError):
        continue
    op = (src[1] if src and src[0].startswith("@") else src[0] if src else "").split(".")[0]
    if op in W:
        flops += W[op] * n; by[op] = by.get(op, 0) + n
print("kernel:", name)
print("executed FP32 flops:", flops, {k: v for k, v in sorted(by.items())})
