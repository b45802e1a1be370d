"""-m gpu: BASELINE.json's full size (C4: 10 M-triangle displaced mesh + 10 k spheres, 1920 x 1080,
depth 16) through the C ABI, checked with size-independent properties plus an oracle subset.

* closest-hit ids / t / u / v of a 200 k-ray subset (camera + bounce rays) bit-exact vs the oracle
* any-hit  <=>  closest-hit found something, for every ray of a 4 M-ray batch
* the result of a ray does not depend on its position in the batch (dynamic ray fetch)
* tmax := t_hit rejects the hit and everything behind it (miss); tmax := next float after t_hit
  returns the same primitive
* frames are independent: render(frames 0..3) == render(0..1) accumulated with render(2..3), bit for bit,
  and every pixel carries its sample count
"""
import os

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi, scenes
from tests import util_scenes as U

pytestmark = pytest.mark.gpu
W, H = 1920, 1080


@pytest.fixture(scope="module")
def c4():
    api = capi.load()
    g = capi.Context(api)
    n = scenes.MESH_N["10M"]
    mesh = scenes.displaced_mesh(n)
    info = scenes.bench_scene(g, n, 10000, W, H, patchwork=True, mesh=mesh)
    return g, mesh, info


def camera_and_bounce_rays(g, n_rays, seed):
    fov_y = scenes.fov_x_to_y(45.0, W, H)
    s2c = scenes.perspective_sample_to_camera(fov_y, np.float32(W) / np.float32(H))
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(**{k: scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up")}))
    px = np.random.RandomState(seed).randint(0, W * H, n_rays // 2).astype(np.uint32)
    cam = scenes.camera_rays(s2c, c2w, W, H, 0, px)
    h = g.trace_closest(cam)
    hit = h["prim"] != capi.MISS
    p = np.stack([cam[k][hit] + h["t"][hit] * cam[d][hit] for k, d in (("ox", "dx"), ("oy", "dy"), ("oz", "dz"))], -1)
    b = U.random_rays(len(p), seed + 1)
    b["ox"], b["oy"], b["oz"] = p.T.astype(np.float32)
    return np.concatenate([cam, b])


def test_c4_trace_properties(c4):
    g, _, info = c4
    assert g.stats()["n_prims"] == info["n_tris"] + 10000 + 2
    rays = camera_and_bounce_rays(g, 4_000_000, 21)
    h = g.trace_closest(rays)
    hit = h["prim"] != capi.MISS
    assert 0.3 < hit.mean() < 1.0
    assert np.array_equal(g.trace_any(rays) != 0, hit)
    perm = np.random.RandomState(5).permutation(len(rays))
    hp = g.trace_closest(rays[perm])
    for f in ("prim", "t", "u", "v"):
        assert np.array_equal(hp[f].view(np.uint32), h[f][perm].view(np.uint32)), f
    r2 = rays[hit].copy()
    r2["tmax"] = h["t"][hit]
    assert (g.trace_closest(r2)["prim"] == capi.MISS).all()
    r2["tmax"] = np.nextafter(h["t"][hit], np.float32(np.inf))
    h2 = g.trace_closest(r2)
    assert np.array_equal(h2["prim"], h["prim"][hit]) and np.array_equal(h2["t"].view(np.uint32), h["t"][hit].view(np.uint32))


def test_c4_trace_subset_matches_oracle(c4, oracle_api):
    g, mesh, _ = c4
    o = pyoracle.OracleContext(oracle_api)
    scenes.bench_scene(o, scenes.MESH_N["10M"], 10000, W, H, patchwork=True, mesh=mesh)
    rays = camera_and_bounce_rays(g, 200_000, 33)
    want, got = o.trace_closest(rays), g.trace_closest(rays)
    assert np.array_equal(got["prim"], want["prim"])
    hit = want["prim"] != capi.MISS
    for f in ("t", "u", "v"):
        assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32)), f
    assert np.array_equal(g.trace_any(rays), o.trace_any(rays))


def test_c4_render_frames_are_independent(c4):
    g, _, _ = c4
    full, _, _ = g.render(W, H, 16, 0, 4)
    assert (full[..., 3] == 4).all() and np.isfinite(full).all()
    acc, _, _ = g.render(W, H, 16, 0, 2)
    g.render_into(capi.RenderParams(W, H, 16, 2, 2, 1, capi.RENDER_HOST_BUFFERS, 0), acc)
    assert np.array_equal(acc.view(np.uint32), full.view(np.uint32))
    st = g.stats()
    assert st["rays_closest"] >= 2 * W * H and st["rays_any"] > 0
    # energy sanity: the image is lit (rectangle light above the terrain) and not saturated
    lum = full[..., :3].mean() / 4
    assert 1e-3 < lum < 10.0
