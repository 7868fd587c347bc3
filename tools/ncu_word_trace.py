#!/usr/bin/env python3
"""One word of the long-document kernel's loop from an ncu source page (--page source --csv --print-source sass):
the instructions executed most often, in order, each with its share of the warp's time in issue-cycle units
(samples / samples-per-issue), the running sum, and its main stall reasons.
usage: tools/ncu_word_trace.py source.csv [first_n]"""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
first_n = int(sys.argv[2]) if len(sys.argv) > 2 else 140
hdr = rows[1]; col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
body = [r for r in rows[2:] if len(r) >= len(hdr)]
cnt = Counter(r[col['Instructions Executed']] for r in body if (r[col['Instructions Executed']] or '0') != '0')
hot_n = max(cnt, key=lambda k: int(k) * cnt[k])
hot = [r for r in body if r[col['Instructions Executed']] == hot_n]
sel = sum(int(r[col['stall_selected']] or 0) for r in hot); tot = sum(int(r[col['# Samples']] or 0) for r in hot)
unit = sel / len(hot)
print(f"hot loop: {len(hot)} instructions x {hot_n} executions, {tot} samples, {tot / unit:.0f} cycles per trip")
agg = Counter()
for r in hot:
    for st in stalls: agg[st[6:]] += int(r[col[st]] or 0)
print("stall mix:", ", ".join(f"{k} {v / tot:.2f}" for k, v in agg.most_common(7)))
cum = 0
for r in hot[:first_n]:
    s = int(r[col['# Samples']] or 0); cum += s
    top = sorted(((int(r[col[st]] or 0), st[6:]) for st in stalls), reverse=True)[:2]
    print(r[col['Address']][-4:], f"{s / unit:6.1f} {cum / unit:7.1f}", r[col['Source']][:60].ljust(60), " ".join(f"{k}:{v / unit:.1f}" for v, k in top if v > 1.5 * unit))
