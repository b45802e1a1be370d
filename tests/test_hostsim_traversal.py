"""BVH build + traversal LOGIC of the CUDA kernels, run on the CPU through tests/hostsim (the same
PB_HD function bodies compiled by g++), against the oracle's brute force and its own binary BVH.
Bit-exact primitive ids and t are required."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi
from tests import util_scenes as U


def hs_build(lib, pos, idx, sph):
    lib.hs_create.restype = C.c_void_p
    lib.hs_create.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
    pos = np.ascontiguousarray(pos, np.float32)
    idx = np.ascontiguousarray(idx, np.uint32)
    sph = np.ascontiguousarray(sph, np.float32)
    M = np.eye(4, dtype=np.float32)[:3].copy()
    h = lib.hs_create(pos.ctypes.data, len(pos), idx.ctypes.data, len(idx), M.ctypes.data, sph.ctypes.data, len(sph))
    return C.c_void_p(h)


def hs_closest(lib, h, rays):
    hits = np.empty(len(rays), capi.HIT_DTYPE)
    lib.hs_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.hs_trace_closest(h, rays.ctypes.data, len(rays), hits.ctypes.data)
    return hits


def hs_any(lib, h, rays):
    occ = np.empty(len(rays), np.uint8)
    lib.hs_trace_any.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.hs_trace_any(h, rays.ctypes.data, len(rays), occ.ctypes.data)
    return occ


@pytest.mark.parametrize("n_tri,n_sph,aa,seed", [(1, 0, False, 1), (0, 1, False, 2), (2, 0, False, 3), (7, 3, False, 4),
                                                 (500, 20, False, 5), (3000, 0, True, 6), (20000, 200, False, 7)])
def test_closest_and_any_match_oracle(hostsim_lib, oracle_api, n_tri, n_sph, aa, seed):
    pos, idx, sph = U.random_soup(n_tri, n_sph, seed, axis_aligned=aa)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    h = hs_build(hostsim_lib, pos, idx, sph)
    n_rays = 4000 if n_tri > 5000 else 20000
    rays = U.random_rays(n_rays, seed + 100, from_points=pos if aa else None)
    rays["tmax"][::3] = 1.5  # bounded rays too
    want = o.trace_closest_brute(rays) if n_tri <= 5000 else o.trace_closest(rays)
    got = hs_closest(hostsim_lib, h, rays)
    assert np.array_equal(got["prim"], want["prim"])
    hit = want["prim"] != capi.MISS
    assert hit.sum() > 0 or n_tri + n_sph < 3
    for f in ("t", "u", "v"):
        assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32))
    # the oracle's own BVH agrees with its brute force as well (tree independence)
    assert np.array_equal(o.trace_closest(rays)["prim"], want["prim"])
    occ_want = o.trace_any_brute(rays) if n_tri <= 5000 else o.trace_any(rays)
    assert np.array_equal(hs_any(hostsim_lib, h, rays), occ_want)
    assert np.array_equal(occ_want != 0, hit)
    hostsim_lib.hs_destroy(h)


def test_empty_scene(hostsim_lib):
    h = hs_build(hostsim_lib, np.zeros((0, 3), np.float32), np.zeros((0, 3), np.uint32), np.zeros((0, 3, 4), np.float32))
    rays = U.random_rays(100, 1)
    assert (hs_closest(hostsim_lib, h, rays)["prim"] == capi.MISS).all()
    assert (hs_any(hostsim_lib, h, rays) == 0).all()


@pytest.mark.parametrize("leaf_max,bottom_max", [(1, 0), (1, 8), (2, 8), (3, 8), (3, 24), (2, 16)])
def test_tree_shape_knobs_do_not_change_results(hostsim_lib, oracle_api, leaf_max, bottom_max):
    """PB_LEAF_MAX / PB_BOTTOM_MAX (primitives per leaf slot, size of the subtrees collapsed into one bottom node)
    change the 8-wide tree, never the hits: the intersection contract is tree independent."""
    pos, idx, sph = U.random_soup(4000, 60, 11)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    rays = U.random_rays(6000, 111)
    rays["tmax"][::4] = 1.2
    want = o.trace_closest_brute(rays)
    hostsim_lib.hs_set_leaf_max.argtypes = [C.c_uint32]
    hostsim_lib.hs_set_bottom_max.argtypes = [C.c_uint32]
    hostsim_lib.hs_set_leaf_max(leaf_max)
    hostsim_lib.hs_set_bottom_max(bottom_max)
    try:
        h = hs_build(hostsim_lib, pos, idx, sph)
        got = hs_closest(hostsim_lib, h, rays)
        assert np.array_equal(got["prim"], want["prim"])
        hit = want["prim"] != capi.MISS
        assert np.array_equal(got["t"][hit].view(np.uint32), want["t"][hit].view(np.uint32))
        assert np.array_equal(hs_any(hostsim_lib, h, rays) != 0, hit)
        hostsim_lib.hs_destroy(h)
    finally:
        hostsim_lib.hs_set_leaf_max(2)
        hostsim_lib.hs_set_bottom_max(8)


@pytest.mark.parametrize("flags", ["-DPB_COLD_SMEM=0 -DPB_TRI_TEST_FIRST=0 -DPB_SPHERE_TEST_FIRST=0",
                                   "-DPB_COLD_SMEM=1 -DPB_TRI_TEST_FIRST=1 -DPB_SPHERE_TEST_FIRST=1"])
def test_compile_time_switches_do_not_change_results(oracle_api, flags, tmp_path):
    """PB_COLD_SMEM (cold ray state behind the shared-memory stack), PB_TRI_TEST_FIRST / PB_SPHERE_TEST_FIRST (primitive test
    before the slab of its padded box) change instruction order and register use, never the accepted
    set: the kernels' bodies built with the switches off and with all of them on return the oracle's brute-force hits bit for bit."""
    import os
    import subprocess
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hostsim")
    so = str(tmp_path / "libhostsim_variant.so")
    fma = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", *fma, "-shared", "-Wno-unknown-pragmas",
                           *flags.split(), "-o", so, os.path.join(d, "hostsim.cpp")])
    lib = C.CDLL(so)
    pos, idx, sph = U.random_soup(3000, 80, 21)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    rays = U.random_rays(8000, 121)
    rays["tmax"][::3] = 1.5
    want = o.trace_closest_brute(rays)
    h = hs_build(lib, pos, idx, sph)
    got = hs_closest(lib, h, rays)
    assert np.array_equal(got["prim"], want["prim"])
    hit = want["prim"] != capi.MISS
    assert hit.sum() > 1000
    for f in ("t", "u", "v"):
        assert np.array_equal(got[f][hit].view(np.uint32), want[f][hit].view(np.uint32))
    assert np.array_equal(hs_any(lib, h, rays), o.trace_any_brute(rays))
    lib.hs_destroy(h)
