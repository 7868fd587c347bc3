#!/usr/bin/env python
"""Turn an .ncu-rep (captured on the B200 box, read here with `ncu -i`) into the short text
summary committed under profiles/:  python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt

Per profiled launch: duration, DRAM bytes, FP64-pipe and issue utilisation, occupancy limiters,
warp-stall mix, and - from the source page - the executed-instruction mix of the hottest loop.
"""
import collections
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__warps_eligible.avg.per_cycle_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]
STALLS = ["long_scoreboard", "wait", "math_pipe_throttle", "not_selected", "short_scoreboard", "selected", "dispatch_stall",
          "branch_resolving", "no_instruction", "lg_throttle", "mio_throttle", "barrier"]


def ncu(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main(rep):
    rows = ncu(rep, "raw")
    hdr, units = rows[0], rows[1]
    print(f"# {rep}")
    for r in rows[2:]:
        rec = dict(zip(hdr, r))
        print(f"\n## launch {rec.get('ID')}: {rec.get('Kernel Name', '')[:100]}")
        for k in RAW_KEYS:
            if k in rec:
                print(f"{k:75s} {rec[k]:>18s} {units[hdr.index(k)]}")
        tot = 0.0
        parts = []
        for s in STALLS:
            k = f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"
            if k in rec:
                try:
                    v = float(rec[k])
                except ValueError:
                    continue
                parts.append((v, s))
                tot += v
        if tot > 0:
            print("warp stalls per issued instruction: " + ", ".join(f"{s} {v:.2f}" for v, s in sorted(parts, reverse=True)[:6]))
    src = ncu(rep, "source")
    # the source page concatenates launches; each block starts with a "Kernel Name" row
    blocks, cur = [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            blocks.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None:
            cur["rows"].append(r)
    for b in blocks:
        h = b["hdr"]
        if not h or "Instructions Executed" not in h:
            continue
        ia, isrc, isamp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
        data = [r for r in b["rows"] if len(r) > ia and r[ia].isdigit()]
        if not data:
            continue
        cnt = collections.Counter(int(r[ia]) for r in data)
        top_count = max(cnt.items(), key=lambda kv: kv[0] * kv[1])[0]
        hot = [r for r in data if int(r[ia]) >= 0.9 * top_count]
        mix = collections.Counter()
        for r in hot:
            t = r[isrc].split()
            mix[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += 1
        total = sum(int(r[ia]) for r in data)
        print(f"\n## source page: {b['name'][:90]}")
        print(f"SASS instructions {len(data)}, executed warp-instructions {total}; hottest loop: {len(hot)} instructions "
              f"x {top_count} trips = {100.0 * len(hot) * top_count / total:.0f}% of all executed")
        print("hottest loop mix: " + ", ".join(f"{op} {n}" for op, n in mix.most_common(12)))
        smp = sum(int(r[isamp]) for r in hot)
        alls = sum(int(r[isamp]) for r in data)
        print(f"stall samples in the hottest loop: {smp} of {alls}")


if __name__ == "__main__":
    main(sys.argv[1])
