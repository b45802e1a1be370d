#!/usr/bin/env python
"""raycast_bench.py — BASELINE.json configs[2] ("C3"): ray-cast microbench against the procedural
displaced mesh (+ spheres): coherent primary rays and incoherent cosine-weighted bounce rays through
pb_trace_closest / pb_trace_any with DEVICE buffers; reports Mrays/s, algorithmic bytes per ray
(counted pass) and the HBM-roofline fraction.  `--check N` compares N rays of each set with the test
oracle (bit-exact ids and t).  Not the headline bench (that is bench.py); one JSON line per ray set."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def bounce_rays(rays, hits, seed=1, target=0):
    """One cosine-sampled bounce ray from every primary hit (geometric normal from the hit triangle is not
    needed for a ray SET: the hemisphere is oriented around +Y flipped against the incoming ray).
    `target` > number of hits: further bounce rays (new directions) from the same hit points until the set
    holds `target` rays (C3 asks for as many incoherent rays as primary rays; only ~1/3 of those hit)."""
    from pupiloptixlab_b200 import capi
    hit = np.flatnonzero(hits["prim"] != capi.MISS)
    if target > len(hit) > 0:
        hit = np.resize(hit, target)                 # cyclic repetition of the hit list
    r, h = rays[hit], hits[hit]
    n = len(r)
    rng = np.random.RandomState(seed)
    u1, u2 = rng.rand(n).astype(np.float32), rng.rand(n).astype(np.float32)
    st, phi = np.sqrt(u1), 2 * np.pi * u2
    lx, lz, ly = st * np.cos(phi), st * np.sin(phi), np.sqrt(np.maximum(0, 1 - u1))
    out = np.zeros(n, capi.RAY_DTYPE)
    out["ox"] = r["ox"] + h["t"] * r["dx"]; out["oy"] = r["oy"] + h["t"] * r["dy"]; out["oz"] = r["oz"] + h["t"] * r["dz"]
    out["dx"], out["dy"], out["dz"] = lx, ly, lz
    out["tmin"], out["tmax"] = 1e-3, 1e16
    return out


def morton_order(rays, lo=-1.0, hi=1.0):
    def q(v):
        return np.clip(((v - lo) / (hi - lo) * 1024).astype(np.int64), 0, 1023)
    def spread(x):
        x = (x | (x << 16)) & 0x030000FF
        x = (x | (x << 8)) & 0x0300F00F
        x = (x | (x << 4)) & 0x030C30C3
        x = (x | (x << 2)) & 0x09249249
        return x
    k = (spread(q(rays["ox"])) << 2) | (spread(q(rays["oy"])) << 1) | spread(q(rays["oz"]))
    return np.argsort(k, kind="stable")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mesh", default="10M", help="1M | 10M | 100M, or a comma list (primary rays are generated once)")
    ap.add_argument("--rays", type=int, default=20_000_000, help="primary rays (C3: 100000000)")
    ap.add_argument("--incoherent", type=int, default=0, help="size of the incoherent set (0 = one bounce ray per primary hit; C3: 100000000)")
    ap.add_argument("--orders", action="store_true", help="also time the incoherent set in primary-pixel order and binned by direction octant")
    ap.add_argument("--no-sorted", action="store_true", help="skip the origin-Morton-sorted variant of the incoherent set")
    ap.add_argument("--spheres", type=int, default=0)
    ap.add_argument("--check", type=int, default=0, help="rays of each set compared with the oracle's BVH traversal (bit-exact ids and t)")
    ap.add_argument("--brute", type=int, default=0, help="rays of each set compared with the oracle's brute-force loop over all primitives")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from pupiloptixlab_b200 import capi, scenes
    api = capi.load()
    side = int(np.sqrt(args.rays))
    fov_y = scenes.fov_x_to_y(45.0, side, side)
    s2c = scenes.perspective_sample_to_camera(fov_y, 1.0)
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(**{k: scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up")}))
    prim = scenes.camera_rays(s2c, c2w, side, side, 0)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev); torch.cuda.set_stream(stream)
    for mesh_name in args.mesh.split(","):
        one_mesh(args, mesh_name, prim, api, capi, scenes, torch, dev, stream, peak)


def one_mesh(args, mesh_name, prim, api, capi, scenes, torch, dev, stream, peak):
    g = capi.Context(api, 0)
    n = scenes.MESH_N[mesh_name]
    t0 = time.time()
    mesh = scenes.displaced_mesh(n)
    t1 = time.time()
    info = scenes.bench_scene(g, n, args.spheres, 1920, 1080, patchwork=False, mesh=mesh)
    st0 = g.stats()
    print(json.dumps({"mesh": mesh_name, "n_tris": info["n_tris"], "mesh_gen_s": round(t1 - t0, 1), "upload_commit_s": round(time.time() - t1, 1),
                      "bvh_build_ms": st0["build_ms"], "bvh_nodes": st0["n_nodes"],
                      "build_mtris_per_s": info["n_tris"] / st0["build_ms"] / 1e3 if st0["build_ms"] else None}), flush=True)

    def run(name, rays, any_hit=False):
        d_r = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).to(dev)
        d_h = torch.empty((len(rays), 4), dtype=torch.float32, device=dev) if not any_hit else torch.empty(len(rays), dtype=torch.uint8, device=dev)
        fn = g.trace_any_dev if any_hit else g.trace_closest_dev
        for _ in range(3):
            fn(d_r.data_ptr(), len(rays), d_h.data_ptr(), stream.cuda_stream)
        ms = []
        for _ in range(args.reps):
            fn(d_r.data_ptr(), len(rays), d_h.data_ptr(), stream.cuda_stream)
            ms.append(g.stats()["last_trace_ms"])
        t = float(np.median(ms))
        cnt_fn = g.api._trace_any_counted if any_hit else g.api._trace_closest_counted
        g._ck(cnt_fn(g.h, d_r.data_ptr(), len(rays), d_h.data_ptr(), 0, stream.cuda_stream), "counted")
        st = g.stats()
        b = (32 + (1 if any_hit else 16)) + (80.0 * st["node_visits"] + 48.0 * st["prim_tests_tri"] + 32.0 * st["prim_tests_sphere"]) / len(rays)
        out = {"set": name, "kernel": "k_trace<any>" if any_hit else "k_trace<closest>", "n_rays": len(rays), "mesh": mesh_name, "n_tris": info["n_tris"],
               "spheres": args.spheres, "ms": t, "mrays_per_s": len(rays) / t / 1e3, "bytes_per_ray": b,
               "nodes_per_ray": st["node_visits"] / len(rays), "tris_per_ray": st["prim_tests_tri"] / len(rays),
               "roofline": {"bound": "hbm", "achieved": len(rays) * b / (t * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": len(rays) * b / (t * 1e-3) / 1e9 / peak}}
        print(json.dumps(out), flush=True)
        res = d_h.cpu().numpy()
        del d_r, d_h
        return res.view(capi.HIT_DTYPE).reshape(-1) if not any_hit else res

    hp = run("coherent primary", prim)
    inc = bounce_rays(prim, hp, target=args.incoherent)
    if args.orders:   # diagnostic: the same rays in the order a wavefront would hold them, and binned by direction octant
        run("incoherent bounce (primary-pixel order)", inc)
        octant = (inc["dx"] < 0).astype(np.int8) * 4 + (inc["dy"] < 0).astype(np.int8) * 2 + (inc["dz"] < 0).astype(np.int8)
        run("incoherent bounce (primary-pixel order, binned by direction octant)", inc[np.argsort(octant, kind="stable")])
        del octant
    rng = np.random.RandomState(5)
    inc = inc[rng.permutation(len(inc))]            # queue order of a wavefront after compaction: unordered
    hi = run("incoherent bounce (unordered)", inc)
    occ = run("incoherent bounce (unordered), any-hit", inc, any_hit=True)
    # size-independent property: "some accepted hit exists" == "the closest hit exists", for every ray
    print(json.dumps({"check": "any-hit == (closest-hit found)", "mesh": mesh_name, "rays": len(inc),
                      "mismatches": int(((hi["prim"] != capi.MISS) != (occ != 0)).sum())}), flush=True)
    del occ
    if not args.no_sorted:
        inc_s = inc[morton_order(inc)]
        run("incoherent bounce (origin-Morton sorted)", inc_s)
        del inc_s
    if args.check:
        from oracle import pyoracle
        o = pyoracle.OracleContext(pyoracle.load())
        scenes.bench_scene(o, n, args.spheres, 1920, 1080, patchwork=False, mesh=mesh)
        bvh = o.import_cwbvh(g)       # the committed 8-wide BVH, read by the oracle (structure + leaf records verified)
        for name, rays, got in (("coherent", prim, hp), ("incoherent", inc, hi)):
            sel = np.random.RandomState(9).randint(0, len(rays), args.check)
            t_o = time.time()
            want = o.trace_closest(rays[sel])
            t_o = time.time() - t_o
            bad = int((want["prim"] != got["prim"][sel]).sum())
            hit = want["prim"] != capi.MISS
            bad_t = int((want["t"][hit].view(np.uint32) != got["t"][sel][hit].view(np.uint32)).sum())
            # algorithmic bytes from the ORACLE's ordered, t-culled traversal of that BVH (a 10^6-ray subset)
            sub = rays[sel[:1_000_000]]
            cw, cnt = o.trace_cwbvh(sub)
            b_orc = 48.0 + (80.0 * cnt["nodes"] + 48.0 * cnt["tris"] + 32.0 * cnt["spheres"]) / len(sub)
            print(json.dumps({"check": name, "mesh": mesh_name, "rays": args.check, "id_mismatches": bad, "t_mismatches": bad_t, "hit_fraction": float(hit.mean()),
                              "oracle_mrays_per_s": args.check / t_o / 1e6,
                              "oracle_ordered_traversal": {"rays": len(sub), "nodes_per_ray": cnt["nodes"] / len(sub), "tris_per_ray": cnt["tris"] / len(sub),
                                                           "spheres_per_ray": cnt["spheres"] / len(sub), "bytes_per_ray": b_orc,
                                                           "id_mismatches_vs_own_bvh": int((cw["prim"] != want["prim"][:len(sub)]).sum())},
                              "bvh": bvh}), flush=True)
        if args.brute:      # ground truth: O(N) loop over every primitive, a small subset
            for name, rays, got in (("coherent", prim, hp), ("incoherent", inc, hi)):
                sel = np.random.RandomState(10).randint(0, len(rays), args.brute)
                want = o.trace_closest_brute(rays[sel])
                print(json.dumps({"check": name + " vs brute force", "mesh": mesh_name, "rays": args.brute,
                                  "id_mismatches": int((want["prim"] != got["prim"][sel]).sum())}), flush=True)
        del o
    del g


if __name__ == "__main__":
    main()
