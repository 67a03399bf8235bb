#!/usr/bin/env python
"""Region-level breakdown of an .ncu-rep source page: contiguous SASS runs with equal execution count.
usage: python profiles/ncu_regions.py rep.ncu-rep [min_pct]"""
import csv, subprocess, sys
rep = sys.argv[1]; minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-count", "1"], capture_output=True, text=True).stdout   # first kernel of the report
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, isrc, ie, it, isamp = (hdr.index(k) for k in ("Address", "Source", "Instructions Executed", "Avg. Threads Executed", "# Samples"))
data = []
for r in rows[2:]:
    try: data.append((int(r[ia], 16), r[isrc], int(r[ie]), float(r[it]), int(r[isamp])))
    except (ValueError, IndexError): pass
base = data[0][0]; tot = sum(d[2] for d in data); tots = sum(d[4] for d in data)
print(f"total warp-instructions {tot:.3e}, samples {tots}, mean active lanes {sum(d[2]*d[3] for d in data)/tot:.2f}")
start = 0
for i in range(1, len(data) + 1):
    if i == len(data) or abs(data[i][2] - data[start][2]) > 0.02 * max(data[start][2], 1):
        e = sum(d[2] for d in data[start:i]); s = sum(d[4] for d in data[start:i])
        if e > tot * minpct / 100 or s > tots * minpct / 100:
            ops = {}
            for d in data[start:i]:
                op = d[1].split()[0] if not d[1].startswith("@") else d[1].split()[1]
                op = op.split(".")[0]; ops[op] = ops.get(op, 0) + 1
            top = " ".join(f"{k}:{v}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:5])
            print(f"off {data[start][0]-base:#7x} n={i-start:4d} count={data[start][2]/1e6:8.2f}M inst%={100*e/tot:5.1f} samp%={100*s/tots:5.1f} lanes={data[start][3]:4.1f} | {top}")
        start = i
