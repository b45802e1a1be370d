"""-m gpu: closest-hit / any-hit through the C ABI of libpupil_b200.so (CUDA, sm_100a) against the
oracle.  Primitive ids, t, u, v must be bit-exact (0 mismatches)."""
import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi, scenes
from tests import util_scenes as U

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    return capi.load()


def _cmp(got, want):
    assert np.array_equal(got["prim"], want["prim"]), f"{(got['prim'] != want['prim']).sum()} id mismatches"
    hit = want["prim"] != capi.MISS
    for f in ("t", "u", "v"):
        assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32)), f
    return hit


@pytest.mark.parametrize("n_tri,n_sph,aa,seed", [(0, 0, False, 0), (1, 0, False, 1), (0, 1, False, 2), (7, 3, False, 4),
                                                 (500, 20, False, 5), (3000, 0, True, 6), (200000, 2000, False, 7)])
def test_trace_matches_oracle(api, oracle_api, n_tri, n_sph, aa, seed):
    pos, idx, sph = U.random_soup(n_tri, n_sph, seed, axis_aligned=aa)
    o = pyoracle.OracleContext(oracle_api)
    g = capi.Context(api)
    U.oracle_scene(o, pos, idx, sph)
    U.oracle_scene(g, pos, idx, sph)
    rays = U.random_rays(200000, seed + 100, from_points=pos if aa else None)
    rays["tmax"][::3] = 1.5
    brute = n_tri <= 3000
    sub = rays[:20000] if brute else rays
    want = o.trace_closest_brute(sub) if brute else o.trace_closest(sub)
    got = g.trace_closest(sub)
    hit = _cmp(got, want)
    occ = g.trace_any(sub)
    assert np.array_equal(occ != 0, hit)
    # counted (instrumented) build returns the same hits and non-zero visit counters
    got2 = g.trace_closest(sub, counted=True)
    _cmp(got2, want)
    st = g.stats()
    if n_tri + n_sph:
        assert st["node_visits"] > 0 and st["n_nodes"] > 0
    # empty ray batch
    assert len(g.trace_closest(rays[:0])) == 0


def test_mesh_1m_coherent_and_incoherent(api, oracle_api):
    """C3-shaped: camera rays + cosine bounce rays against the 1 M-triangle displaced mesh + spheres."""
    n = scenes.MESH_N["1M"]
    mesh = scenes.displaced_mesh(n)
    o = pyoracle.OracleContext(oracle_api)
    g = capi.Context(api)
    for ctx in (o, g):
        info = scenes.bench_scene(ctx, n, 1000, 1920, 1080, patchwork=False, mesh=mesh)
    fov_y = scenes.fov_x_to_y(45.0, 1920, 1080)
    s2c = scenes.perspective_sample_to_camera(fov_y, np.float32(1920) / np.float32(1080))
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(**{k: scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up")}))
    px = np.random.RandomState(3).randint(0, 1920 * 1080, 300000).astype(np.uint32)
    rays = scenes.camera_rays(s2c, c2w, 1920, 1080, 0, px)
    want = o.trace_closest(rays)
    got = g.trace_closest(rays)
    hit = _cmp(got, want)
    assert hit.mean() > 0.3
    # incoherent: random directions from the hit points
    p = np.stack([rays[k][hit] + got["t"][hit] * rays[d][hit] for k, d in (("ox", "dx"), ("oy", "dy"), ("oz", "dz"))], -1)
    r2 = U.random_rays(len(p), 11)
    r2["ox"], r2["oy"], r2["oz"] = p.T.astype(np.float32)
    _cmp(g.trace_closest(r2), o.trace_closest(r2))
    assert np.array_equal(g.trace_any(r2), o.trace_any(r2))
    st = g.stats()
    assert st["n_prims"] == info["n_tris"] + 1000 + 2


def test_counted_traversal_against_oracle_counts(api, oracle_api):
    """B_ray comes from the ORACLE's ordered, t-culled traversal of the committed BVH (pb_export_bvh ->
    oracle/orc_cwbvh.h).  The export passes the oracle's structural checks, the oracle's traversal of the product's tree
    returns the product's hits bit for bit, and the kernel's own counters (PB_*_COUNTED: octant order, batched primitive
    tests, stale stack entries) are compared with the oracle's: the kernel never visits fewer nodes than the ordered
    traversal, and the excess stays bounded (the bench reports both; roofline.achieved uses the oracle's)."""
    n = scenes.MESH_N["1M"]
    mesh = scenes.displaced_mesh(n)
    o = pyoracle.OracleContext(oracle_api)
    g = capi.Context(api)
    for ctx in (o, g):
        scenes.bench_scene(ctx, n, 2000, 1920, 1080, patchwork=True, mesh=mesh)
    info = o.import_cwbvh(g)
    assert info["n_prims"] == g.stats()["n_prims"] and info["n_nodes"] == g.stats()["n_nodes"] and info["n_leaf_slots"] > 0
    fov_y = scenes.fov_x_to_y(45.0, 1920, 1080)
    s2c = scenes.perspective_sample_to_camera(fov_y, np.float32(1920) / np.float32(1080))
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(**{k: scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up")}))
    px = np.random.RandomState(8).randint(0, 1920 * 1080, 100000).astype(np.uint32)
    cam = scenes.camera_rays(s2c, c2w, 1920, 1080, 0, px)
    h0 = g.trace_closest(cam)
    hit = h0["prim"] != capi.MISS
    p = np.stack([cam[k][hit] + h0["t"][hit] * cam[d][hit] for k, d in (("ox", "dx"), ("oy", "dy"), ("oz", "dz"))], -1)
    bounce = U.random_rays(len(p), 12)
    bounce["ox"], bounce["oy"], bounce["oz"] = p.T.astype(np.float32)
    for name, rays in (("camera", cam), ("bounce", bounce)):
        want, oc = o.trace_cwbvh(rays)
        got = g.trace_closest(rays, counted=True)
        st = g.stats()
        _cmp(got, want)
        kc = {"nodes": st["node_visits"], "tris": st["prim_tests_tri"], "spheres": st["prim_tests_sphere"]}
        alg = lambda c: 48.0 + (80 * c["nodes"] + 48 * c["tris"] + 32 * c["spheres"]) / len(rays)
        print(f"{name}: oracle ordered traversal {oc} = {alg(oc):.0f} B/ray; kernel counters {kc} = {alg(kc):.0f} B/ray")
        assert oc["nodes"] <= kc["nodes"] <= 2.0 * oc["nodes"]
        assert alg(oc) <= alg(kc) * 1.02 and alg(kc) <= 2.0 * alg(oc)
        occ, oa = o.trace_cwbvh(rays, any_hit=True)
        assert np.array_equal(g.trace_any(rays, counted=True) != 0, occ != 0)
