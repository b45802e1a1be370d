"""-m gpu: the C++ host layer end to end — headless System + B200PTPass (libpupil_host.so) loads a
mitsuba-style XML, uploads it through the C ABI, renders N frames (1 frame = 1 spp) on cuda:0 and hands
back the "final result" / "albedo" / "normal" buffers; compared with the oracle rendering the same XML."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi, hostlib

pytestmark = pytest.mark.gpu
DATA = os.path.join(os.path.dirname(__file__), "data")


@pytest.mark.parametrize("xml,frames", [("cornell_box.xml", 4), ("cornell_materials.xml", 3)])
def test_headless_pass_matches_oracle(oracle_api, xml, frames):
    path = os.path.join(DATA, xml)
    fin, alb, nrm = hostlib.run_path_tracer(path, frames)
    sc = hostlib.Scene(path=path)
    i = sc.info()
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(o)
    so, ao, no = o.render(i.film_w, i.film_h, i.max_depth, 0, frames)
    mean = so[..., :3] / so[..., 3:4]
    assert (fin[..., 3] == 1).all()
    err = np.abs(fin[..., :3] - mean) / (np.abs(mean) + 1e-3)
    assert (err.max(-1) <= 1e-3).mean() >= 0.998
    assert (np.abs(alb - ao).max(-1) <= 1e-3).mean() >= 0.998
    assert (np.abs(nrm - no).max(-1) <= 1e-3).mean() >= 0.998


def test_xml_upload_into_the_product_and_trace(oracle_api):
    sc = hostlib.Scene(path=os.path.join(DATA, "cornell_materials.xml"))
    g = capi.Context(capi.load())
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(g)
    sc.upload(o)
    from tests import util_scenes as U
    rays = U.random_rays(50000, 3, extent=0.9)
    rays["oy"] += 1.0
    hg, ho = g.trace_closest(rays), o.trace_closest(rays)
    assert np.array_equal(hg["prim"], ho["prim"]) and np.array_equal(hg["t"].view(np.uint32), ho["t"].view(np.uint32))


def test_headless_pass_with_image_files(oracle_api, tmp_path):
    """XML + PNG texture + EXR environment map (tests/test_image_io.py's scene) through System + B200PTPass on the GPU,
    saved as .exr / tone-mapped .png, against the oracle rendering the same loaded scene."""
    from tests import test_image_io as T
    from pupiloptixlab_b200 import scenes
    os.makedirs(tmp_path / "tex")
    T.write_png(tmp_path / "tex" / "albedo.png", np.random.RandomState(7).randint(30, 256, (8, 12, 3)), 2)
    open(tmp_path / "sky.exr", "wb").write(T.exr_file(scenes.procedural_envmap(32, 16)[..., :3], 3))
    path = str(tmp_path / "scene.xml")
    open(path, "w").write(T.XML_TEXTURED)
    fin, alb, nrm = hostlib.run_path_tracer(path, 6)
    sc = hostlib.Scene(path=path)
    o = pyoracle.OracleContext(oracle_api)
    sc.upload(o)
    so, ao, no = o.render(64, 48, 5, 0, 6)
    mean = so[..., :3] / so[..., 3:4]
    err = np.abs(fin[..., :3] - mean) / (np.abs(mean) + 1e-3)
    assert (err.max(-1) <= 1e-3).mean() >= 0.995
    # save + reload (BitmapTexture::Save flips the rows), preview through pb_tonemap
    hostlib.save_image(str(tmp_path / "out.exr"), fin)
    np.testing.assert_array_equal(hostlib.load_image(str(tmp_path / "out.exr"))[::-1, :, :3], fin[..., :3])
    g = capi.Context(capi.load())
    hostlib.save_image(str(tmp_path / "out.png"), g.tonemap(fin))
    assert hostlib.load_image(str(tmp_path / "out.png")).shape == (48, 64, 4)
