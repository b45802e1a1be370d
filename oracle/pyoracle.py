"""ORACLE — test infrastructure, not product code.

ctypes loader for the CPU oracle (oracle/liboracle.so, or oracle/_ref/liboracle_ref.so which also
contains the reference's own headers).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from pupiloptixlab_b200 import capi

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "liboracle_ref.so")


def build(ref: bool | None = None) -> None:
    """Compile the oracle (and oracle/_ref when the reference tree is present)."""
    subprocess.check_call(["make", "-s", "-C", HERE])
    if ref is None:
        ref = os.path.isdir("/root/reference/framework")
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _extend(api: capi.Api) -> capi.Api:
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    fpp = C.POINTER(C.c_float)
    api._f("has_reference", i32, [])
    api._f("set_backend", i32, [vp, i32])
    api._f("max_threads", i32, [])
    api._f("set_threads", None, [i32])
    api._f("trace_closest_brute", i32, [vp, vp, u64, vp])
    api._f("trace_any_brute", i32, [vp, vp, u64, vp])
    api._f("rng_stream", i32, [vp, u32, u32, u32, C.POINTER(u32), fpp])
    api._f("bsdf_table", i32, [vp, u32, fpp, C.POINTER(u32), u32, fpp])
    api._f("emitter_table", i32, [vp, u32, fpp, fpp, fpp, u32, fpp])
    api._f("render_window", i32, [vp, C.POINTER(capi.RenderParams), C.POINTER(capi.RenderTargets), u32, u32, u32, u32, u32])
    api._f("import_cwbvh", i32, [vp, vp, u64, vp, u64])
    api._f("cwbvh_leaf_slots", u64, [vp])
    api._f("trace_cwbvh", i32, [vp, vp, u64, vp, i32, C.POINTER(u64)])
    return api


_cache = {}


def load(prefer_ref: bool = True) -> capi.Api:
    """Load the oracle; prefers the _ref build (superset) when it exists."""
    key = bool(prefer_ref and os.path.exists(REF_LIB))
    if key not in _cache:
        path = REF_LIB if key else PORT_LIB
        if not os.path.exists(path):
            build(ref=False)
        _cache[key] = _extend(capi.Api(path, "orc_"))
    return _cache[key]


class OracleContext(capi.Context):
    """capi.Context plus the oracle-only probes."""

    def __init__(self, api: capi.Api | None = None, backend: int = 0):
        super().__init__(api or load())
        self.set_backend(backend)

    def has_reference(self) -> bool:
        return bool(self.api._has_reference())

    def set_backend(self, backend: int):
        self._ck(self.api._set_backend(self.h, backend), "set_backend")

    def trace_closest_brute(self, rays):
        rays = np.ascontiguousarray(rays, capi.RAY_DTYPE)
        hits = np.empty(rays.shape[0], capi.HIT_DTYPE)
        self._ck(self.api._trace_closest_brute(self.h, rays.ctypes.data, rays.shape[0], hits.ctypes.data), "trace_closest_brute")
        return hits

    def trace_any_brute(self, rays):
        rays = np.ascontiguousarray(rays, capi.RAY_DTYPE)
        occ = np.empty(rays.shape[0], np.uint8)
        self._ck(self.api._trace_any_brute(self.h, rays.ctypes.data, rays.shape[0], occ.ctypes.data), "trace_any_brute")
        return occ

    def render_window(self, width, height, max_depth, first_sample, sample_count, x0, y0, w, h, sample_stride=1, step=1, counted=False):
        """w x h pixels (x0 + i * step, y0 + j * step) of a width x height film (full-film seeds): (sum, albedo, normal,
        test), window-sized.  counted: also run every ray through the imported product BVH (import_cwbvh) and leave the
        visit counts of that ordered traversal in stats()."""
        p = capi.RenderParams(width, height, max_depth, first_sample, sample_count, sample_stride,
                              capi.RENDER_CLEAR | (capi.RENDER_COUNTED if counted else 0), 0)
        acc = np.zeros((h, w, 4), np.float32)
        alb, nrm = np.zeros((h, w, 3), np.float32), np.zeros((h, w, 3), np.float32)
        tst = np.zeros((h, w), np.float32)
        t = capi.RenderTargets(acc.ctypes.data, alb.ctypes.data, nrm.ctypes.data, tst.ctypes.data)
        self._ck(self.api._render_window(self.h, C.byref(p), C.byref(t), x0, y0, w, h, step), "render_window")
        return acc, alb, nrm, tst

    def import_cwbvh(self, product_ctx) -> dict:
        """Export the committed 8-wide BVH of a product context (pb_export_bvh) and import it here; the scene of both
        contexts must be the same.  Returns sizes (nodes, prims, leaf slots)."""
        api = product_ctx.api
        nn, npr = C.c_uint64(), C.c_uint64()
        product_ctx._ck(api._export_bvh(product_ctx.h, None, C.byref(nn), None, C.byref(npr)), "export_bvh")
        nodes = np.zeros(max(1, nn.value) * 80, np.uint8)
        prims = np.zeros(max(1, npr.value) * 12, np.float32)
        product_ctx._ck(api._export_bvh(product_ctx.h, nodes.ctypes.data, C.byref(nn), prims.ctypes.data, C.byref(npr)), "export_bvh")
        self._ck(self.api._import_cwbvh(self.h, nodes.ctypes.data, nn.value, prims.ctypes.data, npr.value), "import_cwbvh")
        return {"n_nodes": nn.value, "n_prims": npr.value, "n_leaf_slots": int(self.api._cwbvh_leaf_slots(self.h))}

    def trace_cwbvh(self, rays, any_hit=False):
        """Ordered, t-culled traversal of the imported BVH: (hits | occluded, {nodes, tris, spheres})."""
        rays = np.ascontiguousarray(rays, capi.RAY_DTYPE)
        out = np.empty(rays.shape[0], np.uint8 if any_hit else capi.HIT_DTYPE)
        cnt = (C.c_uint64 * 3)()
        self._ck(self.api._trace_cwbvh(self.h, rays.ctypes.data, rays.shape[0], out.ctypes.data, int(any_hit), cnt), "trace_cwbvh")
        return out, {"nodes": int(cnt[0]), "tris": int(cnt[1]), "spheres": int(cnt[2])}

    def rng_stream(self, v0: int, v1: int, n: int):
        seed = C.c_uint32()
        out = np.empty(n, np.float32)
        self._ck(self.api._rng_stream(self.h, v0, v1, n, C.byref(seed), capi._fp(out)), "rng_stream")
        return seed.value, out

    def bsdf_table(self, mat: int, wo: np.ndarray, seeds: np.ndarray) -> np.ndarray:
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        seeds = np.ascontiguousarray(seeds, np.uint32)
        out = np.empty((wo.shape[0], 16), np.float32)
        self._ck(self.api._bsdf_table(self.h, mat, capi._fp(wo), seeds.ctypes.data_as(C.POINTER(C.c_uint32)),
                                      wo.shape[0], capi._fp(out)), "bsdf_table")
        return out

    def emitter_table(self, ei: int, hp, hn, xi) -> np.ndarray:
        hp = np.ascontiguousarray(hp, np.float32).reshape(-1, 3)
        hn = np.ascontiguousarray(hn, np.float32).reshape(-1, 3)
        xi = np.ascontiguousarray(xi, np.float32).reshape(-1, 2)
        out = np.empty((hp.shape[0], 16), np.float32)
        self._ck(self.api._emitter_table(self.h, ei, capi._fp(hp), capi._fp(hn), capi._fp(xi), hp.shape[0], capi._fp(out)), "emitter_table")
        return out
