"""-m gpu: BASELINE.json's full size (C4: 10 M-triangle displaced mesh + 10 k spheres, 1920 x 1080,
depth 16) through the C ABI, checked with size-independent properties plus an oracle subset.

* closest-hit ids / t / u / v of a 200 k-ray subset (camera + bounce rays) bit-exact vs the oracle
* any-hit  <=>  closest-hit found something, for every ray of a 4 M-ray batch
* the result of a ray does not depend on its position in the batch (dynamic ray fetch)
* tmax := t_hit rejects the hit and everything behind it (miss); tmax := next float after t_hit
  returns the same primitive
* frames are independent: render(frames 0..3) == render(0..1) accumulated with render(2..3), bit for bit,
  and every pixel carries its sample count
* PIXELS of the headline workload: the full 1920 x 1080 film at 1 and at 2 spp, depth 16, against the oracle on the
  same seeds (radiance, albedo / normal AOVs, the "test" draw), and a 480 x 270 window of the C5 film (3840 x 2160)
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


@pytest.fixture(scope="module")
def c4_oracle(c4, oracle_api):
    _, mesh, _ = c4
    o = pyoracle.OracleContext(oracle_api)
    scenes.bench_scene(o, scenes.MESH_N["10M"], 10000, W, H, patchwork=True, mesh=mesh)
    return o


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


def test_c4_trace_subset_matches_oracle(c4, c4_oracle):
    g, mesh, _ = c4
    o = c4_oracle
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


REL_TOL = 1e-3


def rel_ok_fraction(a, b):
    err = np.abs(a - b) / (np.abs(b) + 1e-3)
    return float((err.max(axis=-1) <= REL_TOL).mean())


def tonemapped_mean(img):
    x = img.astype(np.float64)
    return float((x / (1.0 + x)).mean())


# (first frame, spp, depth, required fraction of pixels within REL_TOL)
C4_PIXEL_CASES = [(0, 1, 2, 0.999), (3, 2, 3, 0.999), (0, 1, 16, 0.95), (5, 2, 16, 0.90)]


@pytest.mark.parametrize("first,spp,depth,need", C4_PIXEL_CASES)
def test_c4_pixels_match_oracle(c4, c4_oracle, first, spp, depth, need):
    """north_star: "a 1080p, 10M-triangle pt+MIS render whose pixels agree with the reference integrator within
    tolerance".  Whole 1920 x 1080 film, identical per-pixel seeds Init(4, pixel, frame); the only scene with interpolated
    vertex normals + uvs on a 10 M-triangle mesh, the hashed seven-material patchwork and 10 k material-cycling spheres.

    * depth 2 and 3 (primary hit + emission, NEE with MIS, BSDF sample, emitter hit with MIS — every material, every
      shading branch): the bound of the Cornell cases, |a-b| / (|b| + 1e-3) <= 1e-3 on >= 99.9 % of the pixels.
    * depth 16 (the benchmark's depth): a path diverges wherever a libdevice-vs-glibc ulp of sinf / cosf flips a branch
      OR is amplified by reflections off the 5-20 mm spheres (direction error x ~2 d / r per bounce).  The floor for two
      EXACT implementations is measured on the CPU alone by scripts/chaos_floor.py (the oracle against itself with
      sinf / cosf nudged by one ulp): 97.0 % of the pixels within 1e-3 at 1 spp, depth 16 (profiles/r02_chaos_floor.md;
      the GPU measures 97.2 %).  Required: >= 95 % at 1 spp (>= 90 % at 2 spp: either sample may diverge), and — because
      diverged paths are still samples of the same estimator — no bias: the tone-mapped film means x / (1 + x) agree to
      1e-3, and the number of closest-hit rays to 2e-3.
    * always: equal sample counts and non-finite counts, albedo / normal AOVs to 99.9 %, the "test" draw bit-exact."""
    g, _, _ = c4
    o = c4_oracle
    so, ao, no, to = o.render_ex(W, H, depth, first, spp)
    sg, ag, ng, tg = g.render_ex(W, H, depth, first, spp)
    assert np.array_equal(sg[..., 3], so[..., 3]) and (sg[..., 3] == spp).all()
    assert np.isfinite(sg).sum() == np.isfinite(so).sum()
    frac = rel_ok_fraction(sg[..., :3], so[..., :3])
    fa, fn = rel_ok_fraction(ag, ao), rel_ok_fraction(ng, no)
    rc_g, rc_o = g.stats()["rays_closest"], o.stats()["rays_closest"]
    tm_g, tm_o = tonemapped_mean(sg[..., :3] / spp), tonemapped_mean(so[..., :3] / spp)
    print(f"C4 depth {depth}, {spp} spp from frame {first}: radiance {frac:.5f}, albedo {fa:.5f}, normal {fn:.5f} of the pixels within "
          f"{REL_TOL}; closest rays {rc_g} vs {rc_o}; tone-mapped film mean {tm_g:.6f} vs {tm_o:.6f}; film mean "
          f"{sg[..., :3].mean() / spp:.6f} vs {so[..., :3].mean() / spp:.6f}")
    assert frac >= need, f"only {frac:.5f} of the pixels within {REL_TOL}"
    assert fa >= 0.999 and fn >= 0.999
    assert np.array_equal(tg.view(np.uint32), to.view(np.uint32))
    assert abs(rc_g - rc_o) <= 2e-3 * rc_o
    assert abs(tm_g - tm_o) <= 1e-3 * tm_o


@pytest.mark.parametrize("depth,spp,need", [(3, 2, 0.999), (16, 1, 0.95)])
def test_c5_film_window_matches_oracle(c4, c4_oracle, depth, spp, need):
    """C5's film (3840 x 2160, same scene): the GPU renders the whole film (8 frames per wave instead of 32, other
    slot tables), the oracle a 480 x 270 window of it with the full-film seeds.  Bounds as in the C4 pixel test."""
    g, _, _ = c4
    o = c4_oracle
    W5, H5 = 3840, 2160
    cam = lambda width, height: [scenes.set_perspective_camera(ctx, scenes.BENCH_CAMERA["fov"], width, height,
                                 scenes.lookat_to_world(*(scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up")))) for ctx in (g, o)]
    cam(W5, H5)      # same scene, the 4K film's sample-to-camera matrix
    try:
        x0, y0, ww, wh = 1712, 700, 480, 270
        so, ao, no, to = o.render_window(W5, H5, depth, 0, spp, x0, y0, ww, wh)
        sg, ag, ng, tg = g.render_ex(W5, H5, depth, 0, spp)
        crop = lambda a: a[y0:y0 + wh, x0:x0 + ww]
        assert (sg[..., 3] == spp).all()
        frac = rel_ok_fraction(crop(sg)[..., :3], so[..., :3])
        print(f"C5 window, depth {depth}, {spp} spp: {frac:.5f} of the pixels within {REL_TOL}")
        assert frac >= need
        assert rel_ok_fraction(crop(ag), ao) >= 0.999 and rel_ok_fraction(crop(ng), no) >= 0.999
        assert np.array_equal(crop(tg).view(np.uint32), to.view(np.uint32))
    finally:
        cam(W, H)


def test_c4_moving_one_sphere_is_a_refit(c4, c4_oracle):
    """VERDICT r1 / SURVEY 8f-3: the reference moves an instance with an IAS *update* build (ias_manager.cpp:116-151), not
    a rebuild.  Moving one of C4's 10 000 spheres: served by the refit (pb_stats.refits), in well under the 70-150 ms a
    rebuild of the 10 M-primitive tree costs (target < 2 ms of library time), captured wave graphs survive (same device
    pointers), and the hits of camera + bounce rays — and the pixels around the sphere — equal the oracle's REBUILT scene."""
    g, _, info = c4
    o = c4_oracle
    obj = 7 + 1234                      # the 7 patchwork meshes come first, then the spheres
    rng = np.random.RandomState(42)     # bench_scene's own generator: recover this sphere's centre / radius
    n_sph = info["n_spheres"]
    ctr = np.stack([rng.uniform(-1, 1, n_sph), rng.uniform(0.05, 0.6, n_sph), rng.uniform(-1, 1, n_sph)], -1)
    rad = rng.uniform(0.005, 0.02, n_sph)
    old = (scenes.translate(*ctr[1234]) @ scenes.scale(rad[1234], rad[1234], rad[1234])).astype(np.float32)
    new = (scenes.translate(0.05, 0.45, 0.9) @ scenes.scale(0.08, 0.08, 0.08)).astype(np.float32)   # in front of the camera, inside the scene box
    g.render(W, H, 16, 0, 1, want_aov=False)                 # captures the wave graph
    try:
        g.update_instance(obj, new[:3])
        st = g.stats()
        print(f"refit of one sphere in the 10 M-primitive scene: {st['build_ms']:.3f} ms, {st['kernel_launches']} launches")
        assert st["refits"] == 1 and st["rebuilds"] == 0
        assert st["build_ms"] < 2.0
        o.update_instance(obj, new[:3])
        rays = camera_and_bounce_rays(g, 200_000, 77)
        want, got = o.trace_closest(rays), g.trace_closest(rays)
        assert np.array_equal(got["prim"], want["prim"])
        hit = want["prim"] != capi.MISS
        for f in ("t", "u", "v"):
            assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32)), f
        sphere_prim = info["n_tris"] + 1234
        assert (got["prim"] == sphere_prim).sum() > 100       # the moved sphere is really in view now
        assert np.array_equal(g.trace_any(rays), o.trace_any(rays))
        # pixels: depth 3 (no chaos amplification yet), whole film through the still-valid wave graph
        so = o.render(W, H, 3, 0, 1)[0]
        sg = g.render(W, H, 3, 0, 1)[0]
        assert rel_ok_fraction(sg[..., :3], so[..., :3]) >= 0.999
    finally:
        g.update_instance(obj, old[:3])
        o.update_instance(obj, old[:3])
    assert g.stats()["refits"] == 2
    rays = camera_and_bounce_rays(g, 100_000, 78)
    assert np.array_equal(g.trace_closest(rays)["prim"], o.trace_closest(rays)["prim"])
