"""Tree-quality experiment on the CPU (tests/hostsim = the kernels' PB_HD bodies compiled by g++): builds the C4
scene (displaced mesh + spheres) with different build knobs and counts node / primitive visits per ray for
primary rays and two generations of cosine-sampled bounce rays.  Results never depend on the tree, so
only the visit counts matter here.

    python scripts/tree_quality.py [--mesh 10M] [--rays 100000] [--leaf 2] [--bottom 8]
"""
from __future__ import annotations

import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pupiloptixlab_b200 import capi, scenes  # noqa: E402
from tests import conftest  # noqa: E402
from tests.test_hostsim_traversal import hs_build, hs_closest  # noqa: E402


def counters(lib, h):
    out = (C.c_ulonglong * 3)()
    lib.hs_counters.argtypes = [C.c_void_p, C.c_void_p]
    lib.hs_counters(h, out)
    return np.array(list(out), np.float64)


def bounce(rays, hits, rng):
    ok = hits["prim"] != 0xFFFFFFFF
    r, t = rays[ok], hits["t"][ok]
    p = np.stack([r["ox"] + t * r["dx"], r["oy"] + t * r["dy"], r["oz"] + t * r["dz"]], -1)
    u1, u2 = rng.uniform(size=len(p)), rng.uniform(size=len(p))
    rr, phi = np.sqrt(u1), 2 * np.pi * u2
    d = np.stack([rr * np.cos(phi), np.sqrt(1 - u1), rr * np.sin(phi)], -1)   # cosine lobe around +Y
    out = np.zeros(len(p), capi.RAY_DTYPE)
    out["ox"], out["oy"], out["oz"] = p.T.astype(np.float32)
    out["dx"], out["dy"], out["dz"] = d.T.astype(np.float32)
    out["tmin"], out["tmax"] = 1e-3, 1e16
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mesh", default="10M")
    ap.add_argument("--rays", type=int, default=100000)
    ap.add_argument("--leaf-sweep", default="", help="comma list of leaf_max values to compare (default: just --leaf)")
    ap.add_argument("--leaf", type=int, default=2)
    ap.add_argument("--bottom", type=int, default=8)
    ap.add_argument("--spheres", type=int, default=10000)
    a = ap.parse_args()
    lib = conftest._build_hostsim() if hasattr(conftest, "_build_hostsim") else None
    if lib is None:
        raise SystemExit("tests/conftest.py has no _build_hostsim()")
    n = scenes.MESH_N[a.mesh]
    pos, _, _, idx = scenes.displaced_mesh(n)
    rng = np.random.RandomState(42)
    ns = a.spheres
    ctr = np.stack([rng.uniform(-1, 1, ns), rng.uniform(0.05, 0.6, ns), rng.uniform(-1, 1, ns)], -1)
    rad = rng.uniform(0.005, 0.02, ns)
    sph = np.zeros((ns, 3, 4), np.float32)
    sph[:, 0, 0] = sph[:, 1, 1] = sph[:, 2, 2] = rad
    sph[:, :, 3] = ctr
    cam = scenes.BENCH_CAMERA
    w, h = 1920, 1080
    s2c = scenes.perspective_sample_to_camera(scenes.fov_x_to_y(cam["fov"], w, h), w / h)
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(cam["origin"], cam["target"], cam["up"]))
    px = np.random.RandomState(1).choice(w * h, a.rays, replace=False)
    primary = scenes.camera_rays(s2c, c2w, w, h, 0, np.sort(px))
    lib.hs_set_bottom_max(a.bottom)
    ref = None
    for leaf in ([int(x) for x in a.leaf_sweep.split(",")] if a.leaf_sweep else [a.leaf]):
        lib.hs_set_leaf_max(leaf)
        t0 = time.time()
        hh = hs_build(lib, pos, idx, sph)
        lib.hs_num_nodes.argtypes = [C.c_void_p]
        lib.hs_levels.argtypes = [C.c_void_p]
        print(f"leaf_max={leaf} bottom_max={a.bottom}: {lib.hs_num_nodes(hh)} nodes, {lib.hs_levels(hh)} levels, build {time.time() - t0:.1f} s", flush=True)
        rays, ids = primary, []
        r2 = np.random.RandomState(7)
        for gen in range(3):
            hits = hs_closest(lib, hh, rays)
            c = counters(lib, hh) / len(rays)
            byts = 48 + 80 * c[0] + 48 * c[1] + 32 * c[2]
            print(f"   gen {gen}: {len(rays)} rays  nodes/ray {c[0]:.2f}  tris/ray {c[1]:.2f}  spheres/ray {c[2]:.2f}  B_ray {byts:.0f}", flush=True)
            ids.append((hits["prim"].copy(), hits["t"].copy()))
            rays = bounce(rays, hits, r2)
        if ref is None:
            ref = ids
        else:
            for (p0, t0_), (p1, t1) in zip(ref, ids):
                assert np.array_equal(p0, p1) and np.array_equal(t0_.view(np.uint32), t1.view(np.uint32)), "tree shape changed a result"
        lib.hs_destroy.argtypes = [C.c_void_p]
        lib.hs_destroy(hh)


if __name__ == "__main__":
    main()
