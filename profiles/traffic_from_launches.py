#!/usr/bin/env python
"""profiles/traffic.json from an ncu launch list (scripts/gpu_call_r2e.sh: `ncu --metrics gpu__time_duration.sum,
dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 900 --csv`).
usage: traffic_from_launches.py profiles/r02_launches_<tag>.csv > profiles/traffic.json   (keeps the old file's "round1" block)
bench.py reads k_extend_dram_bytes_per_launch for `roofline.traffic` / `dram_frac`."""
import csv, json, os, sys
from collections import defaultdict

RAYS_IN_WAVE = 253681509          # closest-hit rays of one 32-spp C4 wave at 1080p, depth 16 (bench.py kernel counters)
LAUNCHES_PER_WAVE = 16

path = sys.argv[1]
rows = defaultdict(dict)
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    rows[int(r["ID"])]["name"] = r["Kernel Name"]
    rows[int(r["ID"])][r["Metric Name"]] = float(r["Metric Value"])
ext = [v for v in rows.values() if "k_extend" in v["name"]]
con = [v for v in rows.values() if "k_connect" in v["name"]]
wave = [v for v in rows.values() if any(k in v["name"] for k in ("k_extend", "k_connect", "k_shade", "k_miss", "k_generate", "k_accumulate", "k_after_shade", "k_prepare", "k_set_frame0", "k_write_aov"))]
dram = lambda v: v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"]
waves = len(ext) / LAUNCHES_PER_WAVE
t_ext = sum(v["gpu__time_duration.sum"] for v in ext)
out = {
    "source": f"{path}: ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 900, "
              f"command: python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu ({len(ext)} k_extend + {len(con)} k_connect launches = {waves:.1f} waves of 32 spp at 1080p, depth 16)",
    "kernel": "k_extend<false>",
    "launches": len(ext),
    "rays_in_wave": RAYS_IN_WAVE,
    "k_extend_dram_bytes_per_launch": round(sum(map(dram, ext)) / len(ext)),
    "dram_bytes_per_ray": round(sum(map(dram, ext)) / (waves * RAYS_IN_WAVE), 1),
    "k_extend_warp_inst_per_ray": round(sum(v["smsp__inst_executed.sum"] for v in ext) / (waves * RAYS_IN_WAVE), 1),
    "warp_inst_source": f"same launch list: smsp__inst_executed.sum {sum(v['smsp__inst_executed.sum'] for v in ext) / 1e9:.2f} G over {waves:.3f} waves x {RAYS_IN_WAVE / 1e6:.1f} M closest-hit rays",
    "k_connect_dram_bytes_per_launch": round(sum(map(dram, con)) / max(len(con), 1)),
    "k_extend_ms_per_wave_under_ncu": round(t_ext / waves / 1e6, 2),
    "k_extend_share_of_launch_time_under_ncu": round(t_ext / sum(v["gpu__time_duration.sum"] for v in wave), 3),
    "note": "per launch like roofline.achieved. DRAM traffic is a fraction of the algorithmic bytes (the upper BVH levels are L2-resident). ncu times are "
            "cold-cache and serialised: the kernel's SHARE of the launch time is what must agree with bench.py's live share_of_step.",
}
old = os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json")
if os.path.exists(old):
    prev = json.load(open(old))
    if "round1" in prev:
        out["round1"] = prev["round1"]
    out["previous"] = {k: prev[k] for k in ("source", "k_extend_dram_bytes_per_launch", "dram_bytes_per_ray", "k_extend_warp_inst_per_ray",
                                             "k_extend_ms_per_wave_under_ncu", "k_extend_share_of_launch_time_under_ncu") if k in prev}
json.dump(out, sys.stdout, indent=1)
print()
