#!/usr/bin/env python
"""dump_sass.py — SASS listings of the hot kernels, straight from the built objects.

    python profiles/dump_sass.py            # writes profiles/sass/<kernel>.sass + profiles/sass/SUMMARY.md

For every kernel: the `cuobjdump -sass -fun` listing (encodings stripped, so that diffs between
rounds stay readable), registers / shared memory / spills from `cuobjdump -res-usage`, and an
opcode histogram with the mnemonics the design leans on called out (128-/256-bit global accesses,
IDP.4A byte conversion, shared-memory stack traffic, warp votes / matches, reductions, local-memory
spills).  Runs here without a GPU."""
from __future__ import annotations

import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "pupiloptixlab_b200", "csrc")
OUT = os.path.join(ROOT, "profiles", "sass")

# (object, regex on the DEMANGLED kernel name, listing name): symbols are looked up in the object, so a changed parameter
# list cannot silently produce an empty listing (round 1's k_accumulate.sass)
KERNELS = [
    ("pb_wavefront.o", r"pb::k_extend<false>", "k_extend"),
    ("pb_wavefront.o", r"pb::k_connect<false>", "k_connect"),
    ("pb_trace.o", r"pb::k_trace<false, false>", "k_trace_closest"),
    ("pb_trace.o", r"pb::k_trace<true, false>", "k_trace_any"),
    ("pb_wavefront.o", r"pb::k_shade<1>", "k_shade_1_diffuse"),
    ("pb_wavefront.o", r"pb::k_shade<3>", "k_shade_3_rough_dielectric"),
    ("pb_wavefront.o", r"pb::k_shade<7>", "k_shade_7_rough_plastic"),
    ("pb_wavefront.o", r"pb::k_generate\(", "k_generate"),
    ("pb_wavefront.o", r"pb::k_accumulate\(", "k_accumulate"),
    ("pb_wavefront.o", r"pb::k_write_aov\(", "k_write_aov"),
    ("pb_wavefront.o", r"pb::k_miss\(", "k_miss"),
    ("pb_build.o", r"pb::k_refit_wide", "k_refit_wide"),
    ("pb_build.o", r"pb::k_update_prims", "k_update_prims"),
]
OPTIONAL = {"k_refit_wide", "k_update_prims"}


def find_symbol(obj_path: str, rx: str) -> str | None:
    """Mangled name of the kernel whose demangled name matches rx."""
    out = subprocess.run(["cuobjdump", "-elf", obj_path], capture_output=True, text=True).stdout
    names = sorted(set(re.findall(r"\.text\.(_Z\w+)", out)))
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    for m, d in zip(names, dem):
        if re.search(rx, d):
            return m
    return None


CALL_OUT = [
    ("LDG.E.128 / LDG.E.ENL2.256 (vector global loads)", r"^LDG\.E\.(128|ENL2\.256)"),
    ("STG.E.128 / STG.E.ENL2.256 (vector global stores)", r"^STG\.E\.(128|ENL2\.256)"),
    ("LDG narrower than 128 bit", r"^LDG\.E(?!\.(128|ENL2))"),
    ("IDP.4A (byte -> float conversion on the FMA-heavy pipe)", r"^IDP\.4A"),
    ("PRMT", r"^PRMT"),
    ("FFMA", r"^FFMA"),
    ("FMNMX / FMNMX3", r"^FMNMX"),
    ("LDS / STS (shared-memory stack)", r"^(LDS|STS)"),
    ("LDL / STL (local memory)", r"^(LDL|STL)"),
    ("VOTE / MATCH / SHFL (warp cooperation)", r"^(VOTE|MATCH|SHFL|WARPSYNC)"),
    ("ATOMG / RED (global atomics)", r"^(ATOMG|RED|ATOM)"),
    ("CCTL / prefetch", r"^CCTL"),
    ("MUFU (special function unit)", r"^MUFU"),
    ("I2F / F2I", r"^(I2F|F2I|I2FP|F2FP)"),
    ("BRA / BSSY / BSYNC (control)", r"^(BRA|BSSY|BSYNC|EXIT|CALL|RET)"),
]


def sh(*cmd):
    return subprocess.run(cmd, capture_output=True, text=True, check=True).stdout


def main():
    os.makedirs(OUT, exist_ok=True)
    lines_md = ["# SASS listings of the hot kernels (sm_100a, nvcc 12.9, `-O3 -fmad=false -lineinfo`)", "",
                "Written by `profiles/dump_sass.py` from the objects `pupiloptixlab_b200/build.py` produces; one",
                "`<kernel>.sass` per kernel next to this file.  Counts are static instructions in the listing.", ""]
    for obj, rx, name in KERNELS:
        path = os.path.join(CSRC, obj)
        if not os.path.exists(path):
            sys.exit(f"{path} missing: run python -m pupiloptixlab_b200.build first")
        mangled = find_symbol(path, rx)
        if mangled is None:
            if name in OPTIONAL:
                continue
            sys.exit(f"no kernel matching {rx!r} in {obj}")
        raw = sh("cuobjdump", "-sass", "-fun", mangled, path)
        res = sh("cuobjdump", "-res-usage", "-fun", mangled, path)
        usage = " ".join(l.strip() for l in res.splitlines() if "REG:" in l)
        ops = collections.Counter()
        listing = []
        for l in raw.splitlines():
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?)\s*;?\s*/\*\s*0x[0-9a-f]+\s*\*/\s*$", l)
            if m:
                ins = m.group(2).rstrip(" ;")
                listing.append(f"/*{m.group(1)}*/  {ins}")
                body = re.sub(r"^@!?U?P\d+\s+", "", ins)
                ops[body.split()[0] if body else "?"] += 1
            elif re.match(r"\s*/\*\s*0x[0-9a-f]+\s*\*/\s*$", l):
                continue                                         # second half of the 128-bit encoding
            elif l.strip().startswith((".L_", "Function", ".headerflags")):
                listing.append(l.strip())
        if not listing:
            sys.exit(f"empty listing for {name} ({mangled})")
        with open(os.path.join(OUT, name + ".sass"), "w") as f:
            f.write(f"// {name}  ({mangled})\n// {usage}\n" + "\n".join(listing) + "\n")
        total = sum(ops.values())
        lines_md += [f"## {name}", "", f"`{usage}` - {total} instructions", "", "| mnemonic group | static count |", "|---|---|"]
        for label, rx in CALL_OUT:
            n = sum(c for op, c in ops.items() if re.match(rx, op))
            if n:
                lines_md.append(f"| {label} | {n} |")
        top = ", ".join(f"{op} {c}" for op, c in ops.most_common(14))
        lines_md += ["", f"Top opcodes: {top}", ""]
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as f:
        f.write("\n".join(lines_md))
    print("\n".join(lines_md))


if __name__ == "__main__":
    main()
