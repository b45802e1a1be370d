#!/usr/bin/env python
"""Per-kernel table of ONE wave from an ncu launch list (the `--metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none` pass of scripts/gpu_call_r2e.sh).
usage: per_kernel_from_launches.py profiles/r02_launches_<tag>.csv [wave index, default 1] > profiles/r02_per_kernel_<tag>.md
ncu times are serialised and cold-cache: the SHARES are what compares with bench.py's live step."""
import csv, sys
from collections import defaultdict

path = sys.argv[1]; wave = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rows = defaultdict(dict)
for r in csv.DictReader([l for l in open(path) if l.startswith('"')]):
    rows[int(r["ID"])]["name"] = r["Kernel Name"]
    rows[int(r["ID"])][r["Metric Name"]] = float(r["Metric Value"])
ids = sorted(rows)
gen = [i for i in ids if "k_generate" in rows[i]["name"]]
s, e = gen[wave], gen[wave + 1] if wave + 1 < len(gen) else ids[-1] + 1
agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
for i in range(s, e):
    v = rows[i]
    n = v["name"].split("(")[0].replace("pb::", "").replace("void ", "")
    if not n.startswith("k_"):
        continue
    a = agg[n]
    a[0] += 1; a[1] += v["gpu__time_duration.sum"] / 1e6; a[2] += v["dram__bytes_read.sum"] / 1e9; a[3] += v["dram__bytes_write.sum"] / 1e9
    a[4] += v["smsp__inst_executed.sum"] / 1e9
T = sum(a[1] for a in agg.values())
print(f"# Kernels of one C4 wave (32 spp at 1080p, depth 16), from `{path.split('/')[-1]}` (wave {wave})\n")
print("Serialised, cold-cache ncu times: compare the shares, not the milliseconds, with `bench.py` (live step of the same build in the bench line).\n")
print("| kernel | launches | ms | share | DRAM read GB | DRAM write GB | DRAM GB/s | warp instr (G) | G instr / ms |")
print("|---|---|---|---|---|---|---|---|---|")
for n in sorted(agg, key=lambda k: -agg[k][1]):
    c, ms, rd, wr, gi = agg[n]
    print(f"| `{n}` | {c} | {ms:.3f} | {100 * ms / T:.1f} % | {rd:.2f} | {wr:.2f} | {(rd + wr) / ms * 1e3:.0f} | {gi:.3f} | {gi / ms:.2f} |")
print(f"| total | {sum(a[0] for a in agg.values())} | {T:.2f} | | {sum(a[2] for a in agg.values()):.1f} | {sum(a[3] for a in agg.values()):.1f} | {sum(a[2] + a[3] for a in agg.values()) / T * 1e3:.0f} | {sum(a[4] for a in agg.values()):.2f} | |")
print("\n## `k_extend` by bounce\n\n| bounce | ms | warp instr (G) | G instr / ms | DRAM GB |\n|---|---|---|---|---|")
b = 0
for i in range(s, e):
    v = rows[i]
    if "k_extend" in v["name"]:
        ms = v["gpu__time_duration.sum"] / 1e6; gi = v["smsp__inst_executed.sum"] / 1e9
        print(f"| {b} | {ms:.3f} | {gi:.3f} | {gi / ms:.2f} | {(v['dram__bytes_read.sum'] + v['dram__bytes_write.sum']) / 1e9:.2f} |"); b += 1
