"""compute-sanitizer is closed on this GPU pool ("runs under it have left GPUs needing a reset"), so the memory-safety
check of the kernels is done where it can be: every PB_HD body the kernels wrap (Morton keys, radix tree, bottom-up boxes,
8-wide collapse + quantisation, traversal with its shared / local stack, refit, uniform-sphere records, slot <-> pixel
maps, the BSDF table body) is compiled with g++ -fsanitize=address,undefined into a second hostsim library and driven
over the same cases as the hostsim tests — in a subprocess with libasan preloaded.  Any out-of-bounds access, misaligned
or overflowing arithmetic in that code aborts the subprocess.  (What this cannot see: races and device-only intrinsics;
those are covered by the result checks of `pytest -m gpu` against the oracle.)"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

DRIVER = r'''
import ctypes as C, sys, numpy as np
sys.path.insert(0, ROOT)
lib = C.CDLL(SO)
from tests import util_scenes as U
from tests.test_hostsim_traversal import hs_build, hs_closest, hs_any
from tests.test_hostsim_refit import move, rot_z
from tests.test_oracle_cwbvh import hs_export
for n_tri, n_sph, aa, seed in [(0, 0, False, 0), (1, 0, False, 1), (0, 1, False, 2), (7, 3, False, 4), (500, 20, False, 5), (3000, 0, True, 6), (20000, 200, False, 7)]:
    pos, idx, sph = U.random_soup(n_tri, n_sph, seed, axis_aligned=aa)
    for leaf_max, bottom in ((2, 8), (3, 16), (1, 4)):
        lib.hs_set_leaf_max(leaf_max); lib.hs_set_bottom_max(bottom)
        h = hs_build(lib, pos, idx, sph)
        rays = U.random_rays(3000, seed + 100, from_points=pos if aa else None)
        rays["tmax"][::3] = 1.5
        a = hs_closest(lib, h, rays); b = hs_any(lib, h, rays)
        assert ((a["prim"] != 0xFFFFFFFF) == (b != 0)).all()
        if n_sph >= 3:
            assert move(lib, h, 1 if n_tri else 0, np.hstack([rot_z(20) @ np.diag([0.3, 0.2, 0.25]), [[0.1], [0.2], [0.3]]])) in (0, 1)
            hs_closest(lib, h, rays)
        hs_export(lib, h)
        lib.hs_destroy(h)
lib.hs_slot_tables.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
for w, hh, K in [(1, 1, 1), (7, 3, 2), (37, 23, 3), (256, 4, 5), (255, 9, 4)]:
    n = w * hh * K
    f, p, s = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    lib.hs_slot_tables(w, hh, K, f.ctypes.data, p.ctypes.data, s.ctypes.data)
    assert sorted(s.tolist()) == list(range(n))
from tests.golden import make_golden as G
lib.hs_bsdf_table.restype = C.c_int
for k, (name, mat) in enumerate(G.golden_materials()):
    wo, seeds = G.bsdf_inputs(300, 100 + k)
    wo, seeds = np.ascontiguousarray(wo), np.ascontiguousarray(seeds)
    out = np.zeros((300, 16), np.float32)
    assert lib.hs_bsdf_table(C.byref(mat), wo.ctypes.data_as(C.POINTER(C.c_float)), seeds.ctypes.data_as(C.POINTER(C.c_uint32)), 300, out.ctypes.data_as(C.POINTER(C.c_float))) == 0
print("kernel bodies under ASan/UBSan: ok")
'''


def test_kernel_bodies_under_asan_ubsan(tmp_path):
    so = str(tmp_path / "libhostsim_asan.so")
    fma = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
    cmd = ["/usr/bin/g++", "-O1", "-g", "-std=c++17", "-fPIC", "-ffp-contract=off", *fma, "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-fno-omit-frame-pointer", "-shared", "-Wno-unknown-pragmas", "-o", so, os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and "asan" in r.stderr.lower() and "cannot find" in r.stderr.lower():
        pytest.skip("libasan is not installed")
    assert r.returncode == 0, r.stderr[-3000:]
    asan = subprocess.run(["/usr/bin/g++", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    ubsan = subprocess.run(["/usr/bin/g++", "-print-file-name=libubsan.so"], capture_output=True, text=True).stdout.strip()
    env = dict(os.environ, LD_PRELOAD=f"{asan}:{ubsan}", ASAN_OPTIONS="detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    code = f"ROOT = {ROOT!r}\nSO = {so!r}\n" + DRIVER
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=1200)
    assert out.returncode == 0, (out.stdout[-500:], out.stderr[-4000:])
    assert "kernel bodies under ASan/UBSan: ok" in out.stdout
