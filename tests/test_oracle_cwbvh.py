"""the algorithmic bytes per ray are defined on "the visit counts of the oracle's ordered, t-culled
traversal of the committed BVH".  oracle/orc_cwbvh.h is that traversal; here it is checked without a GPU on the tree
the product's build code produces when tests/hostsim runs it on the CPU (the same PB_HD bodies as the kernels):

* import validates the export (every primitive id once, leaf records == the oracle's own world-space primitives bit
  for bit, a proper tree whose leaves cover every primitive) and refuses a corrupted one;
* its hits equal the oracle's binary-BVH and brute-force hits bit for bit (the result is tree independent);
* its counts are those of the product traversal's kernel body (octant order, children tested once at the parent,
  deferred entries not re-culled; run lane-serially by hostsim) up to the visiting order: within a factor 1.5 either
  way on triangle soups (neither order is optimal for heavily overlapping boxes), fewer nodes on the displaced mesh.
The same comparison on the GPU's own counters is tests/test_gpu_trace.py::test_counted_traversal_against_oracle_counts.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi
from tests import util_scenes as U
from tests.test_hostsim_traversal import hs_build, hs_closest, hs_any


def hs_export(lib, h):
    lib.hs_num_nodes.argtypes = [C.c_void_p]
    lib.hs_num_prims.argtypes = [C.c_void_p]
    lib.hs_export.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    nn, npr = lib.hs_num_nodes(h), lib.hs_num_prims(h)
    nodes, prims = np.zeros(max(nn, 1) * 80, np.uint8), np.zeros(max(npr, 1) * 12, np.float32)
    lib.hs_export(h, nodes.ctypes.data, prims.ctypes.data)
    return nodes, nn, prims, npr


def hs_counts(lib, h):
    out = (C.c_ulonglong * 3)()
    lib.hs_counters.argtypes = [C.c_void_p, C.c_void_p]
    lib.hs_counters(h, out)
    return {"nodes": int(out[0]), "tris": int(out[1]), "spheres": int(out[2])}


def import_into(o, nodes, nn, prims, npr):
    return o.api._import_cwbvh(o.h, nodes.ctypes.data, nn, prims.ctypes.data, npr)


@pytest.mark.parametrize("n_tri,n_sph,aa,seed", [(1, 0, False, 1), (0, 1, False, 2), (7, 3, False, 4), (500, 20, False, 5),
                                                 (3000, 0, True, 6), (20000, 200, False, 7)])
def test_ordered_traversal_of_the_exported_tree(hostsim_lib, oracle_api, n_tri, n_sph, aa, seed):
    pos, idx, sph = U.random_soup(n_tri, n_sph, seed, axis_aligned=aa)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    h = hs_build(hostsim_lib, pos, idx, sph)
    nodes, nn, prims, npr = hs_export(hostsim_lib, h)
    assert import_into(o, nodes, nn, prims, npr) == 0, o.api._last_error(o.h)
    assert npr == n_tri + n_sph
    rays = U.random_rays(6000, seed + 100, from_points=pos if aa else None)
    rays["tmax"][::3] = 1.5
    want = o.trace_closest_brute(rays) if n_tri <= 5000 else o.trace_closest(rays)
    got, cnt = o.trace_cwbvh(rays)
    assert np.array_equal(got["prim"], want["prim"])
    hit = want["prim"] != capi.MISS
    for f in ("t", "u", "v"):
        assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32))
    occ, cnt_any = o.trace_cwbvh(rays, any_hit=True)
    assert np.array_equal(occ != 0, hit)
    # the product's traversal body on the same tree, same rays (lane-serial on the CPU)
    hs_closest(hostsim_lib, h, rays)
    prod = hs_counts(hostsim_lib, h)
    assert cnt["nodes"] >= 1 and prod["nodes"] >= 1
    alg = lambda c: 80 * c["nodes"] + 48 * c["tris"] + 32 * c["spheres"]
    assert alg(prod) / 1.5 <= alg(cnt) <= alg(prod) * 1.5, (cnt, prod)   # same tree, same rays: only the visiting order differs
    hs_any(hostsim_lib, h, rays)
    prod_any = hs_counts(hostsim_lib, h)
    assert alg(cnt_any) <= alg(cnt) and alg(prod_any) <= alg(prod)
    hostsim_lib.hs_destroy(h)


def test_ordered_traversal_visits_fewer_nodes_on_the_mesh(hostsim_lib, oracle_api):
    """On the benchmark's kind of geometry (displaced grid mesh, camera + bounce rays) the distance-ordered, re-culled
    traversal fetches no more nodes than the kernel body: the difference is the kernel's stale stack entries."""
    from pupiloptixlab_b200 import scenes
    n = 120
    mpos, _, _, midx = scenes._displaced_mesh(n)
    sph = np.zeros((40, 3, 4), np.float32)
    rng = np.random.RandomState(3)
    for k in range(40):
        sph[k, :, :3] = np.eye(3) * rng.uniform(0.01, 0.04)
        sph[k, :, 3] = (rng.uniform(-1, 1), rng.uniform(0.05, 0.5), rng.uniform(-1, 1))
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, mpos, midx, sph)
    h = hs_build(hostsim_lib, mpos, midx, sph)
    assert import_into(o, *hs_export(hostsim_lib, h)) == 0
    fov_y = scenes.fov_x_to_y(45.0, 160, 90)
    s2c = scenes.perspective_sample_to_camera(fov_y, np.float32(160) / np.float32(90))
    c2w = scenes.sensor_to_world(scenes.lookat_to_world(*(scenes.BENCH_CAMERA[k] for k in ("origin", "target", "up"))))
    cam = scenes.camera_rays(s2c, c2w, 160, 90, 0)
    hits = o.trace_closest(cam)
    hit = hits["prim"] != capi.MISS
    p = np.stack([cam[k][hit] + hits["t"][hit] * cam[d][hit] for k, d in (("ox", "dx"), ("oy", "dy"), ("oz", "dz"))], -1)
    b = U.random_rays(len(p), 9)
    b["ox"], b["oy"], b["oz"] = p.T.astype(np.float32)
    for rays in (cam, b):
        got, cnt = o.trace_cwbvh(rays)
        want = hs_closest(hostsim_lib, h, rays)
        assert np.array_equal(got["prim"], want["prim"]) and np.array_equal(got["t"].view(np.uint32), want["t"].view(np.uint32))
        prod = hs_counts(hostsim_lib, h)
        assert cnt["nodes"] <= prod["nodes"], (cnt, prod)
    hostsim_lib.hs_destroy(h)


def test_import_refuses_a_wrong_tree(hostsim_lib, oracle_api):
    pos, idx, sph = U.random_soup(300, 10, 11)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    h = hs_build(hostsim_lib, pos, idx, sph)
    nodes, nn, prims, npr = hs_export(hostsim_lib, h)
    bad = prims.copy(); bad[12 * 5] += 1e-3                       # a moved vertex
    assert import_into(o, nodes, nn, bad, npr) != 0
    bad = prims.copy(); bad[3::12][7] = bad[3::12][8]             # a duplicated primitive id
    assert import_into(o, nodes, nn, bad, npr) != 0
    badn = nodes.copy(); badn[16:20] = 255                        # root child_base outside the node array
    assert import_into(o, badn, nn, prims, npr) != 0
    assert import_into(o, nodes, nn, prims, npr - 1) != 0         # wrong primitive count
    with pytest.raises(capi.ApiError):
        o.trace_cwbvh(U.random_rays(10, 1))                       # nothing imported after the failures
    assert import_into(o, nodes, nn, prims, npr) == 0
    o.trace_cwbvh(U.random_rays(10, 1))
    hostsim_lib.hs_destroy(h)


def test_counted_render_uses_the_imported_tree(hostsim_lib, oracle_api):
    """PB_RENDER_COUNTED on the oracle: every ray of the integrator also walks the imported tree; stats carry the
    ordered-traversal counts, and a disagreement between the two traversals would fail the render."""
    pos, idx, sph = U.random_soup(800, 30, 21)
    o = pyoracle.OracleContext(oracle_api)
    o.scene_begin()
    m = o.add_material(capi.make_material(capi.MAT_DIFFUSE, twosided=True, reflectance=0.6))
    ident = np.eye(4, dtype=np.float32)
    o.add_trimesh(pos, None, None, idx, ident, 0, m)
    for s in sph:
        o.add_sphere(np.vstack([s, [0, 0, 0, 1]]).astype(np.float32), 0, m)
    o.set_env(capi.env_const(0.5, 0.5, 0.5))
    from pupiloptixlab_b200 import scenes
    scenes.set_perspective_camera(o, 60.0, 64, 48, scenes.lookat_to_world((0, 0, 7), (0, 0, 0), (0, 1, 0)))
    o.commit()
    h = hs_build(hostsim_lib, pos, idx, sph)
    assert import_into(o, *hs_export(hostsim_lib, h)) == 0
    plain = o.render(64, 48, 5, 0, 1)[0]
    assert o.stats()["node_visits"] == 0
    counted = o.render_window(64, 48, 5, 0, 1, 0, 0, 64, 48, counted=True)[0]
    st = o.stats()
    assert np.array_equal(plain.view(np.uint32), counted.view(np.uint32))
    assert st["node_visits"] >= st["rays_closest"] > 0 and st["prim_tests_tri"] > 0 and st["node_visits_any"] > 0
    # a strided window: 16 x 12 pixels, every 4th in x and y, are the same pixels of the full film
    sub = o.render_window(64, 48, 5, 0, 1, 1, 2, 15, 11, step=4)[0]
    assert np.array_equal(sub.view(np.uint32), plain[2:2 + 44:4, 1:1 + 60:4].view(np.uint32))
    hostsim_lib.hs_destroy(h)
