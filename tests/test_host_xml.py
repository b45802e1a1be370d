"""The C++ host layer (pupiloptixlab_b200/host -> libpupil_host.so) without a GPU: mitsuba-style XML
loading with the reference's semantics (framework/resource/scene.cpp, shape.cpp, material.cpp,
xml/visitor.h, xml/util_loader.cpp) and the World::LoadScene hand-off through a sink.  The sink used
here is the test oracle's C ABI, so a loaded scene can be rendered on the CPU and compared with the
same scene described directly through the ABI (pupiloptixlab_b200/scenes.py)."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import build as pb_build
from pupiloptixlab_b200 import capi, hostlib, scenes

DATA = os.path.join(os.path.dirname(__file__), "data")


@pytest.fixture(scope="module", autouse=True)
def host_lib():
    pb_build.build_host()
    return hostlib.load()


def tex_rgb(t):
    return tuple(np.float32(x) for x in t.a)


def test_cornell_xml_matches_the_programmatic_scene(oracle_api):
    sc = hostlib.Scene(path=os.path.join(DATA, "cornell_box.xml"))
    i = sc.info()
    assert (i.max_depth, i.film_w, i.film_h) == (8, 256, 256)          # $defaults substituted
    assert (i.n_shapes, i.n_emissive_shapes, i.n_triangles, i.n_spheres, i.has_env) == (8, 1, 2 * 6 + 12 * 2, 0, 0)
    assert sc.warnings() == ["XML Node [sampler] skip", "XML Node [rfilter] skip"]   # unknown tags skipped with their subtree
    s2c, c2w = sc.camera()
    np.testing.assert_allclose(s2c, scenes.perspective_sample_to_camera(scenes.fov_x_to_y(19.5, 256, 256), 1.0), atol=1e-6)
    np.testing.assert_array_equal(c2w, scenes.sensor_to_world(scenes.CORNELL_SENSOR))
    for k, (name, (kind, m)) in enumerate(scenes.CORNELL.items()):
        s = sc.shape(k)
        assert s["type"] == kind
        np.testing.assert_array_equal(s["to_world"], m[:3])
        assert s["material"].type == capi.MAT_DIFFUSE and s["material"].twosided == 1     # <ref id> + twosided wrapper
        assert s["is_emitter"] == (name == "Light")
    assert tex_rgb(sc.shape(4)["material"].reflectance) == (np.float32(0.63), np.float32(0.065), np.float32(0.05))
    assert tex_rgb(sc.shape(7)["radiance"]) == (17, 12, 4)
    # render the uploaded scene with the oracle and compare with the scene built through the ABI
    a = pyoracle.OracleContext(oracle_api)
    sc.upload(a)
    b = pyoracle.OracleContext(oracle_api)
    scenes.cornell_box(b, 256, 256, "diffuse")
    ra, rb = a.render(96, 96, 8, 0, 4)[0], b.render(96, 96, 8, 0, 4)[0]
    err = np.abs(ra - rb)[..., :3] / (np.abs(rb[..., :3]) + 1e-3)
    assert (err.max(-1) <= 1e-3).mean() >= 0.999
    assert a.stats()["n_prims"] == 36 and a.stats()["n_emitters"] == 2


def test_materials_xml_all_seven_bsdfs_and_spheres(oracle_api):
    sc = hostlib.Scene(path=os.path.join(DATA, "cornell_materials.xml"))
    i = sc.info()
    assert (i.n_shapes, i.n_spheres, i.n_emissive_shapes) == (14, 6, 2)
    types = [sc.shape(k)["material"].type for k in range(14)]
    assert set(types) == {1, 2, 3, 4, 5, 6, 7}
    floor = sc.shape(0)["material"]
    assert floor.reflectance.type == capi.TEX_CHECKERBOARD and floor.reflectance.uv_r0[0] == 8 and floor.reflectance.uv_r1[1] == 8
    # sphere: to_world * Translate(center) * Scale(radius)  (shape.cpp:106-125,196-197)
    s = sc.shape(8)
    np.testing.assert_allclose(s["to_world"], [[0.25, 0, 0, -0.55], [0, 0.25, 0, 0.25], [0, 0, 0.25, 0.55]], atol=1e-7)
    a = pyoracle.OracleContext(oracle_api)
    sc.upload(a)
    b = pyoracle.OracleContext(oracle_api)
    scenes.cornell_box(b, 192, 192, "materials")
    ra, rb = a.render(64, 64, 6, 0, 4)[0], b.render(64, 64, 6, 0, 4)[0]
    err = np.abs(ra - rb)[..., :3] / (np.abs(rb[..., :3]) + 1e-3)
    assert (err.max(-1) <= 1e-3).mean() >= 0.995


XML_FEATURES = """<scene version="3.0.0">
  <default name="r" value="0.5"/>
  <default name="rr" value="0.25"/>
  <integrator type="path"/>
  <sensor type="perspective">
    <float name="fov" value="40"/> <string name="fov_axis" value="y"/>
    <transform name="to_world"><lookat origin="0, 1, 5" target="0, 1, 0" up="0, 1, 0"/></transform>
    <film type="hdrfilm"/>
  </sensor>
  <bsdf type="dielectric" id="glass"><string name="int_ior" value="diamond"/><string name="ext_ior" value="water"/></bsdf>
  <bsdf type="conductor" id="gold"><string name="material" value="Au"/></bsdf>
  <shape type="sphere" id="s0"><float name="radius" value="$rr"/><ref id="glass"/>
    <transform name="to_world"><translate x="1" y="2" z="3"/><scale value="2"/><rotate y="1" angle="90"/></transform></shape>
  <shape type="cube" id="c0"><boolean name="flip_normals" value="true"/><ref id="gold"/>
    <transform name="to_world"><matrix value="1 0 0 0 1 0 0 0 1"/></transform></shape>
  <shape type="rectangle" id="r0"><bsdf type="roughplastic"><boolean name="nonlinear" value="true"/>
      <string name="distribution" value="beckmann"/></bsdf></shape>
  <shape type="teapot" id="nope"/>
  <emitter type="constant"><rgb name="radiance" value="$r"/></emitter>
  <emitter type="point"><point name="position" value="0,0,0"/><rgb name="intensity" value="1"/></emitter>
</scene>"""


def test_xml_semantics():
    sc = hostlib.Scene(text=XML_FEATURES)
    i = sc.info()
    assert (i.max_depth, i.film_w, i.film_h) == (1, 768, 576)            # reference defaults (scene.cpp:79-95)
    assert abs(i.fov_y - 40.0) < 1e-6 and i.has_env == capi.ENV_CONSTANT  # fov_axis y: no conversion
    assert i.n_shapes == 3 and any("teapot" in w for w in sc.warnings())
    glass = sc.shape(0)["material"]
    assert glass.type == capi.MAT_DIELECTRIC and glass.int_ior == np.float32(2.419) and glass.ext_ior == np.float32(1.333)
    gold = sc.shape(1)["material"]
    assert gold.type == capi.MAT_CONDUCTOR and tex_rgb(gold.eta) == (np.float32(0.14282), np.float32(0.37414), np.float32(1.43944))
    assert tex_rgb(gold.k) == (np.float32(3.97472), np.float32(2.38066), np.float32(1.59981))
    assert sc.shape(1)["flags"] == capi.SHAPE_FLIP_NORMALS
    # 9-value matrix fills the upper-left 3x3 (util_loader.cpp:137-142)
    np.testing.assert_array_equal(sc.shape(1)["to_world"], np.eye(4, dtype=np.float32)[:3])
    # fixed order scale -> rotate -> translate whatever the XML order, each left-multiplying; then the
    # sphere's own Translate(center) * Scale(radius) on the right
    M = scenes.translate(1, 2, 3) @ scenes.rotate(0, 1, 0, 90) @ scenes.scale(2, 2, 2) @ scenes.scale(0.25, 0.25, 0.25)
    np.testing.assert_allclose(sc.shape(0)["to_world"], M[:3], atol=1e-6)
    rp = sc.shape(2)["material"]
    assert rp.type == capi.MAT_ROUGH_PLASTIC and rp.nonlinear == 1 and rp.int_ior == np.float32(1.49) and rp.twosided == 0
    assert tex_rgb(rp.alpha) == (np.float32(0.1),) * 3 and tex_rgb(rp.reflectance) == (np.float32(0.5),) * 3
    # <lookat> sensor: flipped in the transform loader AND in the sensor case = lookat matrix itself
    _, c2w = sc.camera()
    want = scenes.sensor_to_world(scenes.lookat_to_world((0, 1, 5), (0, 1, 0), (0, 1, 0)))
    np.testing.assert_allclose(c2w, want, atol=1e-6)


def test_xml_errors_are_reported():
    with pytest.raises(hostlib.HostError):
        hostlib.Scene(text="<scene><shape type='cube'></scene>")
    with pytest.raises(hostlib.HostError):
        hostlib.Scene(text="<notscene/>")
    with pytest.raises(hostlib.HostError):
        hostlib.Scene(path="/nonexistent/scene.xml")
    with pytest.raises(hostlib.HostError):
        hostlib.Scene(text="<scene><shape type='cube'><bsdf type='diffuse'><texture type='bitmap' name='reflectance'>"
                           "<string name='filename' value='a.png'/></texture></bsdf></shape></scene>")


def test_obj_loader(tmp_path, oracle_api):
    (tmp_path / "quad.obj").write_text("v -1 0 -1\nv 1 0 -1\nv 1 0 1\nv -1 0 1\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 1 0\n"
                                       "f 1/1/1 2/2/1 3/3/1 4/4/1\n")
    xml = ("<scene><shape type='obj' id='q'><string name='filename' value='quad.obj'/>"
           "<bsdf type='diffuse'/><emitter type='area'><rgb name='radiance' value='1,2,3'/></emitter></shape></scene>")
    sc = hostlib.Scene(text=xml, root_dir=str(tmp_path))
    i = sc.info()
    assert (i.n_shapes, i.n_triangles, i.n_emissive_shapes) == (1, 2, 1)
    assert sc.shape(0)["flags"] == capi.SHAPE_FLIP_TEX_COORDS              # flip_tex_coords defaults to true for obj (shape.cpp:138)
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(o)
    assert o.stats()["n_prims"] == 2 and o.stats()["n_emitters"] == 2
