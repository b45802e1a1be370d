"""pb_update_instance's REFIT (IAS::Update semantics, framework/world/ias_manager.cpp:116-151) on the CPU: tests/hostsim
runs the same PB_HD bodies the kernels wrap (leaf records rewritten by make_prim, refit_wide_node level by level).
After moving / scaling / rotating objects — spheres of both record kinds and the mesh — closest-hit ids, t, u, v and any-hit
flags equal the oracle's REBUILT scene bit for bit, the refitted tree still passes the oracle's structural import of
pb_export_bvh's format, and a move that changes the scene's largest extent is reported as "rebuild" (delta, the padding
of the intersection contract, would change)."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi
from tests import util_scenes as U
from tests.test_hostsim_traversal import hs_build, hs_closest, hs_any
from tests.test_oracle_cwbvh import hs_export, import_into


def move(lib, h, obj, M):
    lib.hs_move_object.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    lib.hs_move_object.restype = C.c_int
    m = np.ascontiguousarray(M, np.float32).reshape(-1)[:12].copy()
    return lib.hs_move_object(h, obj, m.ctypes.data)


def rot_z(deg):
    a = np.deg2rad(deg)
    return np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])


@pytest.mark.parametrize("seed", [3, 4])
def test_refit_equals_the_oracles_rebuilt_scene(hostsim_lib, oracle_api, seed):
    pos, idx, sph = U.random_soup(1500, 40, seed)
    # two far-away anchor spheres (the last two objects, never moved) pin the scene's largest extent, so that the edits
    # below keep delta and are served by the refit
    anchors = np.zeros((2, 3, 4), np.float32)
    for k, c in enumerate((-9.0, 9.0)):
        anchors[k, :, :3] = np.eye(3) * 0.05
        anchors[k, :, 3] = c
    sph = np.concatenate([sph, anchors])
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    h = hs_build(hostsim_lib, pos, idx, sph)
    rng = np.random.RandomState(seed)
    rays = U.random_rays(15000, seed + 50)
    rays["tmax"][::4] = 2.0
    # object 0 is the mesh, objects 1.. are the spheres (even index: s * identity -> "uniform" record, odd: general 3x3)
    edits = [(1, np.hstack([np.eye(3) * 0.3, [[0.5], [-0.4], [1.0]]])),                 # uniform sphere moved + rescaled
             (2, np.hstack([rot_z(30) @ np.diag([0.5, 0.2, 0.3]), [[-1.0], [0.7], [0.2]]])),   # general sphere: rotated ellipsoid
             (3, np.hstack([rot_z(10) @ np.eye(3) * 0.25, [[0.1], [0.1], [-1.2]]])),    # uniform -> rotated (record kind changes)
             (0, np.hstack([rot_z(25) @ np.eye(3) * 0.8, [[0.3], [-0.2], [0.1]]])),     # the whole mesh
             (7, np.hstack([np.eye(3) * 0.15, [[1.5], [1.5], [-1.5]]]))]
    for obj, M in edits:
        rc = move(hostsim_lib, h, obj, M)
        o.update_instance(obj, M)
        assert rc == 0
        want = o.trace_closest(rays)
        got = hs_closest(hostsim_lib, h, rays)
        assert np.array_equal(got["prim"], want["prim"]), f"after moving object {obj}"
        hit = want["prim"] != capi.MISS
        for f in ("t", "u", "v"):
            assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32)), (obj, f)
        assert np.array_equal(hs_any(hostsim_lib, h, rays), o.trace_any(rays))
        # the refitted tree is still a valid export for the oracle's reader (leaf records == the oracle's moved primitives)
        assert import_into(o, *hs_export(hostsim_lib, h)) == 0, o.api._last_error(o.h)
        got_cw, _ = o.trace_cwbvh(rays)
        assert np.array_equal(got_cw["prim"], want["prim"])
    hostsim_lib.hs_destroy(h)


def test_move_that_changes_delta_asks_for_a_rebuild(hostsim_lib):
    pos, idx, sph = U.random_soup(200, 6, 9)
    h = hs_build(hostsim_lib, pos, idx, sph)
    far = np.hstack([np.eye(3) * 0.2, [[50.0], [0.0], [0.0]]])
    assert move(hostsim_lib, h, 2, far) == 1          # the scene's largest extent grows -> delta changes
    hostsim_lib.hs_destroy(h)
