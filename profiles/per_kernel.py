#!/usr/bin/env python
"""per_kernel.py — fold an `ncu --metrics ... --csv` log (one row per launch and metric) into one row per
kernel: launches, total time, share, lanes per instruction (warp execution efficiency), achieved occupancy,
issue-slot utilisation, DRAM and L2 bytes and the bandwidths they imply.  Ratios are time-weighted means.
usage: per_kernel.py gpurun_out/<log>.csv > profiles/<name>.md"""
import csv, collections, re, sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]
ix = {n: i for i, n in enumerate(h)}
launch = collections.OrderedDict()
for r in rows[1:]:
    d = launch.setdefault(r[ix["ID"]], {"kernel": r[ix["Kernel Name"]]})
    v = r[ix["Metric Value"]].replace(",", "")
    try:
        val = float(v)
    except ValueError:
        continue
    unit = r[ix["Metric Unit"]]
    name = r[ix["Metric Name"]]
    if name == "gpu__time_duration.sum":
        val *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)      # -> ms
    elif unit in ("Kbyte", "Mbyte", "Gbyte", "Tbyte"):
        val *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    d[name] = val

def short(k):
    k = re.sub(r"^void ", "", k)
    k = re.sub(r"\(.*$", "", k)
    return k.replace("pb::", "")

agg = collections.OrderedDict()
for d in launch.values():
    a = agg.setdefault(short(d["kernel"]), collections.defaultdict(float))
    t = d.get("gpu__time_duration.sum", 0.0)
    a["n"] += 1; a["ms"] += t
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "smsp__inst_executed.sum"):
        a[k] += d.get(k, 0.0)
    for k in ("smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
              "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"):
        a["w:" + k] += d.get(k, 0.0) * t
tot = sum(a["ms"] for a in agg.values())
print("| kernel | launches | time ms | share | lanes / instr (of 32) | achieved occupancy % | issue active % | L1 / L2 hit % | DRAM GB (r+w) | DRAM GB/s | L2 GB/s |")
print("|---|---|---|---|---|---|---|---|---|---|---|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
    t = a["ms"] or 1e-9
    w = lambda m: a["w:" + m] / t
    dr = a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]
    print(f"| `{k}` | {int(a['n'])} | {a['ms']:.2f} | {100 * a['ms'] / tot:.1f} % | {w('smsp__thread_inst_executed_per_inst_executed.ratio'):.1f} | "
          f"{w('sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | {w('smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{w('l1tex__t_sector_hit_rate.pct'):.0f} / {w('lts__t_sector_hit_rate.pct'):.0f} | {dr / 1e9:.2f} | {dr / 1e6 / t:.0f} | {a['lts__t_bytes.sum'] / 1e6 / t:.0f} |")
