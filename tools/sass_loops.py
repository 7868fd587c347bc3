#!/usr/bin/env python3
"""Lists the loops (backward branches) of one kernel in `cuobjdump -sass` output with their
instruction mix, to check an inner loop's FP64 / other instruction counts before spending GPU time.
Forward branches inside a loop that jump over a block containing a CALL (the out-of-line rare
paths of estep.cu) are followed, so the mix is that of the common path.
usage: cuobjdump -sass build/estep.cu.o | tools/sass_loops.py estep_kernelILi4ELi5E [min_instructions] [-v]"""
import collections
import re
import sys

pat = sys.argv[1]
min_len = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 40
verbose = "-v" in sys.argv
ins = []
on = False
for line in sys.stdin:
    if "Function :" in line:
        on = pat in line
        continue
    if not on:
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
print(f"{len(ins)} instructions")


def opcode(t):
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    return t.split()[0].split(".")[0]


FP64 = ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX")
for i, (a, text) in enumerate(ins):
    m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", text)
    if not m:
        continue
    tgt = int(m.group(1), 16)
    if not (tgt < a and tgt in addr_index):
        continue
    lo, hi = addr_index[tgt], i
    if hi - lo + 1 < min_len:
        continue
    path = []
    k = lo
    while k <= hi:
        a2, t2 = ins[k]
        path.append((a2, t2))
        mb = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t2)
        if mb and k != hi:
            t3 = int(mb.group(1), 16)
            if t3 > a2 and t3 in addr_index and addr_index[t3] <= hi + 1:
                skipped = ins[k + 1:addr_index[t3]]
                if any("CALL" in s for _, s in skipped) or not t2.startswith("@"):
                    k = addr_index[t3]
                    continue
        k += 1
    mix = collections.Counter(opcode(t) for _, t in path)
    f64 = sum(v for kk, v in mix.items() if kk in FP64)
    top = ", ".join(f"{kk} {v}" for kk, v in mix.most_common(16))
    print(f"loop {tgt:#x}..{a:#x}: {hi - lo + 1} instr, common path {len(path)}, FP64 {f64}; {top}")
    if verbose:
        for a2, t2 in path:
            print(f"    {a2:#06x} {t2}")
