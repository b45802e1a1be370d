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
