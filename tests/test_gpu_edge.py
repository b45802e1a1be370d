"""-m gpu: the branches of the hot path the Cornell variants never reach, and the argument-range behaviour of
pb_render, against the oracle on identical seeds (tolerance of tests/test_gpu_render.py).

* edge scene (scenes.edge_scene): constant environment emitter (emitter/env.h:67-86), flip_normals / flip_tex_coords
  (geometry.h:62-75), meshes without normals (cross-product branch, geometry.h:57-60) and without uvs, nonlinear
  plastic / rough plastic, dielectric and rough dielectric with int_ior < ext_ior, non-uniformly scaled spheres
  (one an area emitter);
* the "test" buffer of the pass contract (pt_pass.cpp:144-146, main.cu:101) — an RNG draw, so bit-exact;
* the reference's own fixtures data/static/material_test.xml and mis.xml (copied as tests/data/ref_*.xml) through the
  C++ host layer (System + B200PTPass);
* a scene without emitters, max_depth as the reference's int (-1, 0, 1, 300, > limit), a failed commit.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi, hostlib, scenes

pytestmark = pytest.mark.gpu
REL_TOL, PIXEL_FRACTION = 1e-3, 0.999
DATA = os.path.join(os.path.dirname(__file__), "data")


@pytest.fixture(scope="module")
def api():
    return capi.load()


def rel_ok_fraction(a, b):
    err = np.abs(a - b) / (np.abs(b) + 1e-3)
    return float((err.max(axis=-1) <= REL_TOL).mean())


@pytest.mark.parametrize("w,h,depth,spp,env", [(160, 120, 8, 1, True), (160, 120, 8, 8, True), (96, 72, 5, 4, False), (61, 47, 16, 3, True)])
def test_edge_scene_matches_oracle(api, oracle_api, w, h, depth, spp, env):
    o, g = pyoracle.OracleContext(oracle_api), capi.Context(api)
    scenes.edge_scene(o, w, h, env=env)
    scenes.edge_scene(g, w, h, env=env)
    so, ao, no, to = o.render_ex(w, h, depth, 0, spp)
    sg, ag, ng, tg = g.render_ex(w, h, depth, 0, spp)
    assert np.array_equal(sg[..., 3], so[..., 3]) and (sg[..., 3] == spp).all()
    assert np.isfinite(sg).sum() == np.isfinite(so).sum()
    frac = rel_ok_fraction(sg[..., :3], so[..., :3])
    assert frac >= PIXEL_FRACTION, f"only {frac:.5f} of the pixels within {REL_TOL}"
    assert rel_ok_fraction(ag, ao) >= PIXEL_FRACTION and rel_ok_fraction(ng, no) >= PIXEL_FRACTION
    assert np.array_equal(tg.view(np.uint32), to.view(np.uint32)), "test buffer (main.cu:101) must be bit-exact"
    rc_g, rc_o = g.stats()["rays_closest"], o.stats()["rays_closest"]
    assert abs(rc_g - rc_o) <= 2e-3 * rc_o


def test_edge_scene_every_branch_is_on_screen(api):
    """Guards the scene itself: each special object covers pixels, so the branches really execute on the GPU."""
    g = capi.Context(api)
    w, h = 160, 120
    info = scenes.edge_scene(g, w, h)
    fov_y = scenes.fov_x_to_y(42.0, w, h)
    s2c = scenes.perspective_sample_to_camera(fov_y, np.float32(w) / np.float32(h))
    c2w = scenes.sensor_to_world(scenes.lookat_to_world((0.2, 1.3, 3.4), (0.0, 0.6, 0.0), (0, 1, 0)))
    hits = g.trace_closest(scenes.camera_rays(s2c, c2w, w, h, 0))
    # global prim ids in insertion order: floor 2, back 2, left 2, cube 12, cube2 12, then one id per sphere, light 2
    first = {"floor": 0, "back": 2, "left": 4, "cube": 6, "cube2": 18, "glass": 30, "ellipsoid": 31, "ellipsoid_glass": 32,
             "ellipsoid_light": 33, "light": 34}
    count = {"floor": 2, "back": 2, "left": 2, "cube": 12, "cube2": 12, "light": 2}
    for name, f in first.items():
        if name == "light":
            continue                # the plain rectangle light hangs above the view; it only lights the scene
        n = count.get(name, 1)
        px = int(((hits["prim"] >= f) & (hits["prim"] < f + n)).sum())
        assert px >= (3 if "light" in name else 50), f"{name} covers only {px} pixels"
    assert int((hits["prim"] == capi.MISS).sum()) > 100      # the constant environment is seen directly
    assert set(info["objects"]) == set(first)


def test_scene_without_emitters(api, oracle_api):
    """No area emitter, no environment: the reference would dereference an empty emitter group (emitter.h:110-136);
    here (and in the oracle) paths end at the primary vertex — black film, valid AOVs, no fault."""
    o, g = pyoracle.OracleContext(oracle_api), capi.Context(api)
    w, h = 64, 48
    scenes.edge_scene(o, w, h, emitters=False)
    scenes.edge_scene(g, w, h, emitters=False)
    so, ao, no, to = o.render_ex(w, h, 8, 0, 2)
    sg, ag, ng, tg = g.render_ex(w, h, 8, 0, 2)
    assert (sg[..., :3] == 0).all() and (so[..., :3] == 0).all() and (sg[..., 3] == 2).all()
    assert rel_ok_fraction(ag, ao) >= PIXEL_FRACTION and rel_ok_fraction(ng, no) >= PIXEL_FRACTION
    assert np.array_equal(tg.view(np.uint32), to.view(np.uint32))
    assert g.stats()["n_emitters"] == 0 and g.stats()["rays_any"] == 0
    # the context is still healthy: a lit scene renders afterwards
    scenes.cornell_box(g, 32, 32, "diffuse")
    assert g.render(32, 32, 4, 0, 1)[0][..., :3].max() > 0


@pytest.mark.parametrize("max_depth", [-1, 0, 1, 2, 300])
def test_max_depth_is_the_references_int(api, oracle_api, max_depth):
    """config.max_depth is an int (path_tracer/type.h:12): the loop breaks when ++depth >= max_depth (main.cu:104-106),
    so mitsuba's -1, 0 and 1 all render the primary hit only; 300 exceeds the old 8-bit depth field."""
    o, g = pyoracle.OracleContext(oracle_api), capi.Context(api)
    w = h = 40
    scenes.cornell_box(o, w, h, "materials")
    scenes.cornell_box(g, w, h, "materials")
    md = max_depth & 0xFFFFFFFF
    so = o.render(w, h, md, 0, 2)[0]
    sg = g.render(w, h, md, 0, 2)[0]
    assert rel_ok_fraction(sg[..., :3], so[..., :3]) >= 0.998
    if max_depth <= 1:
        assert g.stats()["rays_any"] == 0 and g.stats()["rays_closest"] == 2 * w * h
        ref = g.render(w, h, 1, 0, 2)[0]
        assert np.array_equal(ref.view(np.uint32), sg.view(np.uint32))


def test_max_depth_above_the_limit_is_refused(api):
    g = capi.Context(api)
    scenes.cornell_box(g, 16, 16, "diffuse")
    with pytest.raises(capi.ApiError, match=r"\(4\)"):       # PB_ERR_LIMIT
        g.render(16, 16, 5000, 0, 1)
    assert g.render(16, 16, 4, 0, 1)[0][..., 3].min() == 1    # and the context still works


def test_failed_commit_leaves_the_context_uncommitted(api):
    """A commit that fails half way must not leave `committed` set over freed device tables."""
    g = capi.Context(api)
    scenes.cornell_box(g, 16, 16, "diffuse")
    g.render(16, 16, 3, 0, 1)
    g.scene_begin()
    m = g.add_material(capi.make_material(capi.MAT_DIFFUSE))
    scenes.add_rectangle(g, np.eye(4, dtype=np.float32), m)
    g.set_env(capi.env_map(np.ones((1, 1, 4), np.float32)))   # 1 x 1 environment map: refused by commit
    with pytest.raises(capi.ApiError):
        g.commit()
    with pytest.raises(capi.ApiError, match=r"\(5\)"):       # PB_ERR_NOT_COMMITTED, not a launch on freed pointers
        g.render(16, 16, 3, 0, 1)
    with pytest.raises(capi.ApiError):
        g.trace_closest(np.zeros(4, capi.RAY_DTYPE))
    scenes.cornell_box(g, 16, 16, "diffuse")
    assert g.render(16, 16, 3, 0, 1)[0][..., :3].max() > 0


@pytest.mark.parametrize("xml,frames", [("ref_material_test.xml", 2), ("ref_mis.xml", 2)])
def test_reference_fixture_scenes_through_b200ptpass(oracle_api, xml, frames):
    """The reference's own scenes (data/static/material_test.xml: nine spheres over a 100 x 100 checkerboard-alpha
    rough-conductor floor, Ag from the IOR table; mis.xml: three sphere lights over four rough-conductor plates) at
    their own film size (1280 x 720) through System + B200PTPass, against the oracle fed by the same loader."""
    path = os.path.join(DATA, xml)
    ses = hostlib.Session(path)
    ses.run(frames)
    fin, alb, nrm, tst, acc = (ses.read(n) for n in ("final result", "albedo", "normal", "test", "pt accum buffer"))
    assert ses.sample_count() == frames
    sc = hostlib.Scene(path=path)
    i = sc.info()
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(o)
    so, ao, no, to = o.render_ex(i.film_w, i.film_h, i.max_depth, 0, frames)
    assert (acc[..., 3] == frames).all() and (fin[..., 3] == 1).all()
    frac = rel_ok_fraction(acc[..., :3], so[..., :3])
    assert frac >= PIXEL_FRACTION, f"{xml}: only {frac:.5f} of the pixels within {REL_TOL}"
    mean = so[..., :3] / so[..., 3:4]
    assert rel_ok_fraction(fin[..., :3], mean) >= PIXEL_FRACTION
    assert rel_ok_fraction(alb, ao) >= PIXEL_FRACTION and rel_ok_fraction(nrm, no) >= PIXEL_FRACTION
    assert np.array_equal(tst.view(np.uint32), to.view(np.uint32))


def test_moving_an_object_through_the_event_path(oracle_api):
    """RenderObject::UpdateTransform -> EWorldEvent::RenderInstanceTransform -> World -> RenderInstanceUpdate -> the
    pass (render_object.cpp:46-49, world.cpp:45-58, pt_pass.cpp:216-218, :40-49): accumulation restarts against the MOVED
    geometry and its rebuilt emitter tables (an emissive sphere is moved), equal to the oracle's edited scene."""
    path = os.path.join(DATA, "ref_mis.xml")
    sc = hostlib.Scene(path=path)
    i = sc.info()
    lights = [k for k in range(i.n_shapes) if sc.shape(k)["is_emitter"]]
    obj = lights[0]
    m = np.eye(4, dtype=np.float32)
    m[:3, :] = sc.shape(obj)["to_world"]
    m[:3, 3] += (0.8, 0.6, -0.5)
    ses = hostlib.Session(path)
    ses.run(3)
    before = ses.read("pt accum buffer")
    ses.move_object(obj, m)
    ses.run(2)
    assert ses.sample_count() == 2           # restarted (pt_pass.cpp:43-44)
    acc = ses.read("pt accum buffer")
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(o)
    o.update_instance(obj, m[:3])
    so = o.render(i.film_w, i.film_h, i.max_depth, 0, 2)[0]
    assert rel_ok_fraction(acc[..., :3], so[..., :3]) >= PIXEL_FRACTION
    assert not np.array_equal(before[..., :3] / 3 * 2, acc[..., :3])
