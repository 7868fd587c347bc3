#!/usr/bin/env python3
"""Prints the hot instructions of a kernel from `ncu -i rep --page source --csv --print-source sass`:
per instruction the stall samples and their main reasons, for the instructions executed most.
usage: tools/ncu_source_hot.py source.csv [min_samples]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
min_s = int(sys.argv[2]) if len(sys.argv) > 2 else 50
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = 0
out = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        n = int(r[col["# Samples"]] or 0)
    except ValueError:
        continue
    tot += n
    top = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:3]
    out.append((r[col["Address"]], r[col["Source"]], n, int(r[col["Instructions Executed"]] or 0), top))
print("total samples", tot)
for addr, src, n, ex, top in out:
    if n >= min_s:
        print(f"{addr[-5:]} {n:7d} {ex:11d}  {src[:70]:70s} " + " ".join(f"{k}:{v}" for v, k in top if v))
