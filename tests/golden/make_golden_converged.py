#!/usr/bin/env python
"""Converged Cornell images from oracle/_ref (backend 1: the reference's own BSDF / emitter / RNG
headers compiled from /root/reference) for the SURVEY 8(c)(ii) protocol: C1 (diffuse, 256 x 256,
depth 8) at 4096 spp and the seven-material variant (128 x 128, depth 8) at 1024 spp.  Dev container
only; about 9 minutes on 8 cores.  Writes tests/golden/converged.npz (per-pixel radiance sums).

    python tests/golden/make_golden_converged.py
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from pupiloptixlab_b200 import scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CONVERGED = [("diffuse", 256, 256, 8, 4096), ("materials", 128, 128, 8, 1024)]


def main():
    api = pyoracle.load(prefer_ref=True)
    if not pyoracle.OracleContext(api).has_reference():
        raise SystemExit("oracle/_ref is not built (needs /root/reference): run `make -C oracle ref` in the dev container")
    out = {}
    for variant, w, h, depth, spp in CONVERGED:
        o = pyoracle.OracleContext(api, backend=1)
        scenes.cornell_box(o, w, h, variant)
        t = time.time()
        s, _, _ = o.render(w, h, depth, 0, spp)
        print(f"{variant} {w}x{h} d{depth} {spp} spp: {time.time() - t:.0f} s", flush=True)
        out[f"{variant}_{w}x{h}_d{depth}_s{spp}_sum"] = s[..., :3].astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "converged.npz"), **out)
    print("wrote converged.npz")


if __name__ == "__main__":
    main()
