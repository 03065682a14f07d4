#!/usr/bin/env python
"""Instruction mix and hottest SASS lines of one kernel from `ncu --page source --csv`.
Usage: ncu -i rep --page source --csv > f.csv; python tools/ncu_source_mix.py f.csv [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
src, ex, samp = ix["Source"], ix["Instructions Executed"], ix["# Samples"]
data = []
for r in rows[2:]:
    try:
        data.append((int(r[ex]), int(r[samp]), r[src].strip()))
    except (ValueError, IndexError):
        pass
tot = sum(n for n, _, _ in data)
tots = sum(s for _, s, _ in data)
print("kernel:", rows[0][1][:100])
print("warp instructions executed:", tot, " stall samples:", tots)
ops = collections.Counter()
for n, s, t in data:
    parts = t.split()
    op = parts[1] if parts and parts[0].startswith("@") else (parts[0] if parts else "?")
    ops[op.split(".")[0]] += n
print("-- instruction mix")
for k, v in ops.most_common(18):
    print(f"  {k:12s} {v:11d} {100 * v / tot:5.1f}%")
print("-- lines with most stall samples")
for n, s, t in sorted(data, key=lambda d: -d[1])[:top]:
    print(f"  {s:6d} samples {n:9d} exec  {t[:90]}")
