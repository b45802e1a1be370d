#!/usr/bin/env python
"""How far can two CORRECT implementations of the path tracer drift apart on the headline scene?

The CUDA path and the CPU oracle evaluate the same fp32 expressions in the same order (-fmad=false / -ffp-contract=off)
and differ only through libdevice's sinf / cosf (1-2 ulp) against glibc's.  This script measures what such a difference
does by itself, on the CPU, with nothing of the product involved: the oracle is compiled a second time with sinf / cosf
results moved by one ulp for a pseudo-random half of the arguments (scripts/chaos_nudge.h), and both oracles render the
C4 scene (BASELINE.json configs[3]) on identical seeds.  At depth 2-3 the images agree to the 1e-3 / 99.9 % bound of the
parity tests; at depth 16 the ulp is amplified by every reflection off a 5-20 mm sphere (a direction error grows by about
2 d / r per bounce, ~ 20-40x here) and ~3 % of the pixels leave the bound although both renders are exact.  The C4 pixel
test (tests/test_gpu_fullsize.py::test_c4_pixels_match_oracle) is calibrated on this floor.

    python scripts/chaos_floor.py [1M|10M] [width height]
"""
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from pupiloptixlab_b200 import capi, scenes  # noqa: E402


def build_nudged(out_dir):
    so = os.path.join(out_dir, "liboracle_nudged.so")
    fma = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++20", "-fPIC", "-fopenmp", "-ffp-contract=off", *fma, "-fvisibility=hidden", "-w",
                           "-include", os.path.join(ROOT, "scripts", "chaos_nudge.h"), "-shared", "-o", so,
                           os.path.join(ROOT, "oracle", "oracle.cpp")])
    return so


def run(mesh_tag="1M", w=960, h=540, depths=(2, 3, 4, 8, 16), n_spheres=10000, out_dir="/tmp"):
    nudged = build_nudged(out_dir)
    pyoracle.build(ref=False)
    n = scenes.MESH_N[mesh_tag]
    mesh = scenes.displaced_mesh(n)
    res = {}
    for name, path in (("plain", pyoracle.PORT_LIB), ("nudged", nudged)):
        o = pyoracle.OracleContext(pyoracle._extend(capi.Api(path, "orc_")))
        scenes.bench_scene(o, n, n_spheres, w, h, patchwork=True, mesh=mesh)
        for d in depths:
            res[(name, d)] = o.render(w, h, d, 0, 1)[0][..., :3]
    rows = []
    for d in depths:
        a, b = res[("nudged", d)], res[("plain", d)]
        err = np.abs(a - b) / (np.abs(b) + 1e-3)
        tm = lambda x: (x / (1 + x)).mean()
        rows.append(dict(depth=d, within_1e3=float((err.max(-1) <= 1e-3).mean()), bit_equal=float((a == b).all(-1).mean()),
                         mean_ratio=float(a.mean() / b.mean()), tonemapped_mean_ratio=float(tm(a) / tm(b))))
    return rows


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "1M"
    w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (960, 540)
    t0 = time.time()
    for r in run(tag, w, h):
        print(r, flush=True)
    print(f"{time.time() - t0:.0f} s")
