"""ctypes view of the C ABI declared in include/pupil_b200.h.

`Api` wraps one shared library that exports that ABI under a symbol prefix.  The product is
`libpupil_b200.so` (prefix ``pb_``, hand-written sm_100a CUDA) and is loaded by `load()`; it raises
if the library is missing — there is no CPU fallback in this package.  (The test oracle exports the
same prototypes under ``orc_`` and re-uses this class from oracle/pyoracle.py.)
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libpupil_b200.so")

PB_OK = 0
TEX_RGB, TEX_BITMAP, TEX_CHECKERBOARD = 0, 1, 2
MAT_DIFFUSE, MAT_DIELECTRIC, MAT_ROUGH_DIELECTRIC, MAT_CONDUCTOR, MAT_ROUGH_CONDUCTOR, MAT_PLASTIC, MAT_ROUGH_PLASTIC = range(1, 8)
SHAPE_FLIP_NORMALS, SHAPE_FLIP_TEX_COORDS = 1, 2
ENV_NONE, ENV_CONSTANT, ENV_MAP = 0, 1, 2
RENDER_HOST_BUFFERS, RENDER_CLEAR, RENDER_COUNTED, RENDER_STAGE_TIMING, RENDER_ASYNC, RENDER_REDUCE = 1, 2, 4, 8, 16, 32
ERR_NCCL = 6
MISS = 0xFFFFFFFF


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("a", C.c_float * 3), ("b", C.c_float * 3),
                ("uv_r0", C.c_float * 4), ("uv_r1", C.c_float * 4),
                ("bitmap", C.POINTER(C.c_float)), ("bitmap_w", C.c_uint32), ("bitmap_h", C.c_uint32),
                ("address_mode", C.c_uint32), ("filter_mode", C.c_uint32)]


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("twosided", C.c_uint32), ("int_ior", C.c_float), ("ext_ior", C.c_float),
                ("nonlinear", C.c_uint32), ("_pad", C.c_uint32),
                ("reflectance", Texture), ("specular_reflectance", Texture), ("specular_transmittance", Texture),
                ("alpha", Texture), ("eta", Texture), ("k", Texture)]


class Env(C.Structure):
    _fields_ = [("type", C.c_uint32), ("radiance", C.c_float * 3), ("scale", C.c_float),
                ("to_world", C.c_float * 12), ("map", Texture)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("max_depth", C.c_uint32),
                ("first_sample", C.c_uint32), ("sample_count", C.c_uint32), ("sample_stride", C.c_uint32),
                ("flags", C.c_uint32), ("_pad", C.c_uint32)]


class RenderTargets(C.Structure):
    _fields_ = [("accum_sum_rgba", C.c_void_p), ("albedo_rgb", C.c_void_p), ("normal_rgb", C.c_void_p), ("test", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("rays_closest", C.c_uint64), ("rays_any", C.c_uint64), ("samples", C.c_uint64),
                ("node_visits", C.c_uint64), ("prim_tests_tri", C.c_uint64), ("prim_tests_sphere", C.c_uint64),
                ("n_prims", C.c_uint64), ("n_nodes", C.c_uint64), ("n_emitters", C.c_uint64),
                ("bvh_depth", C.c_uint32), ("n_objects", C.c_uint32),
                ("build_ms", C.c_float), ("last_render_ms", C.c_float), ("last_trace_ms", C.c_float),
                ("kernel_launches", C.c_uint32),
                ("node_visits_any", C.c_uint64), ("prim_tests_tri_any", C.c_uint64), ("prim_tests_sphere_any", C.c_uint64),
                ("extend_ms", C.c_float), ("connect_ms", C.c_float),
                ("extend_launches", C.c_uint32), ("connect_launches", C.c_uint32),
                ("shade_ms", C.c_float), ("shade_groups", C.c_uint32), ("refits", C.c_uint32), ("rebuilds", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


RAY_DTYPE = np.dtype([("ox", "f4"), ("oy", "f4"), ("oz", "f4"), ("tmin", "f4"),
                      ("dx", "f4"), ("dy", "f4"), ("dz", "f4"), ("tmax", "f4")])
HIT_DTYPE = np.dtype([("t", "f4"), ("u", "f4"), ("v", "f4"), ("prim", "u4")])


def rgb_texture(r, g=None, b=None) -> Texture:
    if g is None:
        g = b = r
    t = Texture()
    t.type = TEX_RGB
    t.a[:] = (r, g, b)
    t.uv_r0[:] = (1, 0, 0, 0)
    t.uv_r1[:] = (0, 1, 0, 0)
    return t


def checker_texture(c_patch1, c_patch2, su=1.0, sv=1.0) -> Texture:
    t = Texture()
    t.type = TEX_CHECKERBOARD
    t.a[:] = c_patch1
    t.b[:] = c_patch2
    t.uv_r0[:] = (su, 0, 0, 0)
    t.uv_r1[:] = (0, sv, 0, 0)
    return t


_KEEPALIVE: list = []
ADDRESS_WRAP, ADDRESS_CLAMP, ADDRESS_MIRROR, ADDRESS_BORDER = 0, 1, 2, 3
FILTER_POINT, FILTER_LINEAR = 0, 1


def bitmap_texture(rgba, address_mode=ADDRESS_WRAP, filter_mode=FILTER_LINEAR, su=1.0, sv=1.0) -> Texture:
    """Bitmap texture from an (h, w, 3|4) float array (util::Texture bitmap, framework/util/texture.h:28-47).
    The array is kept alive on the returned struct; the library copies the texels when the texture is added."""
    a = np.asarray(rgba, np.float32)
    if a.shape[-1] == 3:
        a = np.concatenate([a, np.ones(a.shape[:2] + (1,), np.float32)], -1)
    a = np.ascontiguousarray(a, np.float32)
    t = Texture()
    t.type = TEX_BITMAP
    t.uv_r0[:] = (su, 0, 0, 0)
    t.uv_r1[:] = (0, sv, 0, 0)
    t.bitmap = a.ctypes.data_as(C.POINTER(C.c_float))
    t.bitmap_h, t.bitmap_w = a.shape[0], a.shape[1]
    t.address_mode, t.filter_mode = address_mode, filter_mode
    _KEEPALIVE.append(a)     # ctypes copies the struct by value into Material / Env: the texels must outlive those copies
    if len(_KEEPALIVE) > 64:
        del _KEEPALIVE[0]
    return t


def env_const(r, g, b) -> "Env":
    e = Env()
    e.type = ENV_CONSTANT
    e.radiance[:] = (r, g, b)
    e.scale = 1.0
    e.to_world[:] = (1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0)
    return e


def env_map(rgba, scale=1.0, to_world=None) -> "Env":
    """resource::EnvMap (framework/resource/emitter.h): lat-long radiance bitmap, scale, 3x4 to_world."""
    e = Env()
    e.type = ENV_MAP
    e.scale = scale
    tw = np.eye(4, dtype=np.float32)[:3] if to_world is None else np.asarray(to_world, np.float32).reshape(-1)[:12].reshape(3, 4)
    e.to_world[:] = [float(x) for x in tw.reshape(-1)]
    e.map = bitmap_texture(rgba, ADDRESS_WRAP, FILTER_LINEAR)
    return e


def make_material(mtype, twosided=False, int_ior=1.5046, ext_ior=1.000277, nonlinear=False, reflectance=0.5,
                  specular_reflectance=1.0, specular_transmittance=1.0, alpha=0.1, eta=0.0, k=1.0) -> Material:
    """Defaults follow framework/resource/material.cpp:27-147."""
    def tex(v):
        if isinstance(v, Texture):
            return v
        if np.isscalar(v):
            return rgb_texture(float(v))
        return rgb_texture(*[float(x) for x in v])
    m = Material()
    m.type = mtype
    m.twosided = int(twosided)
    m.int_ior, m.ext_ior = int_ior, ext_ior
    m.nonlinear = int(nonlinear)
    m.reflectance = tex(reflectance)
    m.specular_reflectance = tex(specular_reflectance)
    m.specular_transmittance = tex(specular_transmittance)
    m.alpha = tex(alpha)
    m.eta = tex(eta)
    m.k = tex(k)
    return m


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


class ApiError(RuntimeError):
    pass


class Api:
    """One loaded library exporting the pupil_b200 C ABI under `prefix`."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise ApiError(f"{path} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        f = self._f
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        fpp = C.POINTER(C.c_float)
        f("create", i32, [C.POINTER(vp), i32])
        f("destroy", None, [vp])
        f("last_error", C.c_char_p, [vp])
        f("version", C.c_char_p, [])
        f("scene_begin", i32, [vp])
        f("add_material", i32, [vp, C.POINTER(Material), C.POINTER(u32)])
        f("add_trimesh", i32, [vp, fpp, fpp, fpp, u32, C.POINTER(u32), u32, fpp, u32, u32, C.POINTER(Texture), C.POINTER(u32)])
        f("add_sphere", i32, [vp, fpp, u32, u32, C.POINTER(Texture), C.POINTER(u32)])
        f("set_env", i32, [vp, C.POINTER(Env)])
        f("set_camera", i32, [vp, fpp, fpp])
        f("scene_commit", i32, [vp])
        f("update_instance", i32, [vp, u32, fpp])
        f("trace_closest", i32, [vp, vp, u64, vp, i32, vp])
        f("trace_any", i32, [vp, vp, u64, vp, i32, vp])
        f("render", i32, [vp, C.POINTER(RenderParams), vp, vp, vp, vp])
        f("render_ex", i32, [vp, C.POINTER(RenderParams), C.POINTER(RenderTargets), vp])
        f("resolve", i32, [vp, vp, vp, u32, i32, vp])
        f("get_stats", i32, [vp, C.POINTER(Stats)])
        if self.prefix == "pb_":      # the oracle is synchronous and has no display transform
            f("wait", i32, [vp])
            f("tonemap", i32, [vp, vp, vp, u32, i32, i32, i32, vp])

    def _f(self, name, res, args, optional=False):
        try:
            fn = getattr(self.lib, self.prefix + name)
        except AttributeError:
            if optional:
                return None
            raise
        fn.restype = res
        fn.argtypes = args
        setattr(self, "_" + name, fn)
        return fn

    def version(self) -> str:
        return self._version().decode()


class Context:
    """A scene + renderer instance behind the C ABI (pb_ctx / orc_ctx)."""

    def __init__(self, api: Api, device: int = 0):
        self.api = api
        self.h = C.c_void_p()
        rc = api._create(C.byref(self.h), device)
        if rc != PB_OK:
            msg = api._last_error(self.h).decode() if self.h else "create failed"
            raise ApiError(f"{api.prefix}create failed ({rc}): {msg}")
        self._keep = []

    def close(self):
        if self.h:
            self.api._destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != PB_OK:
            raise ApiError(f"{self.api.prefix}{what} failed ({rc}): {self.api._last_error(self.h).decode()}")

    # ---- scene
    def scene_begin(self):
        self._ck(self.api._scene_begin(self.h), "scene_begin")

    def add_material(self, m: Material) -> int:
        out = C.c_uint32()
        self._ck(self.api._add_material(self.h, C.byref(m), C.byref(out)), "add_material")
        return out.value

    def add_trimesh(self, pos, nrm, uv, idx, to_world, flags, material, area_radiance: Texture | None = None) -> int:
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.uint32).reshape(-1, 3)
        nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 3)
        uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
        tw = np.ascontiguousarray(to_world, np.float32).reshape(-1)[:12].copy()
        out = C.c_uint32()
        ar = C.byref(area_radiance) if area_radiance is not None else None
        self._ck(self.api._add_trimesh(self.h, _fp(pos), _fp(nrm), _fp(uv), pos.shape[0],
                                       idx.ctypes.data_as(C.POINTER(C.c_uint32)), idx.shape[0], _fp(tw), flags, material,
                                       ar, C.byref(out)), "add_trimesh")
        return out.value

    def add_sphere(self, to_world, flags, material, area_radiance: Texture | None = None) -> int:
        tw = np.ascontiguousarray(to_world, np.float32).reshape(-1)[:12].copy()
        out = C.c_uint32()
        ar = C.byref(area_radiance) if area_radiance is not None else None
        self._ck(self.api._add_sphere(self.h, _fp(tw), flags, material, ar, C.byref(out)), "add_sphere")
        return out.value

    def set_env(self, env: Env):
        self._ck(self.api._set_env(self.h, C.byref(env)), "set_env")

    def set_camera(self, sample_to_camera, camera_to_world):
        a = np.ascontiguousarray(sample_to_camera, np.float32).reshape(16).copy()
        b = np.ascontiguousarray(camera_to_world, np.float32).reshape(16).copy()
        self._ck(self.api._set_camera(self.h, _fp(a), _fp(b)), "set_camera")

    def commit(self):
        self._ck(self.api._scene_commit(self.h), "scene_commit")

    def update_instance(self, obj: int, to_world):
        tw = np.ascontiguousarray(to_world, np.float32).reshape(-1)[:12].copy()
        self._ck(self.api._update_instance(self.h, obj, _fp(tw)), "update_instance")

    # ---- ray casting on host numpy arrays (copies are inside the call)
    def trace_closest(self, rays: np.ndarray, counted: bool = False) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.empty(rays.shape[0], HIT_DTYPE)
        fn = self.api._trace_closest_counted if counted else self.api._trace_closest
        self._ck(fn(self.h, rays.ctypes.data, rays.shape[0], hits.ctypes.data, 1, None), "trace_closest")
        return hits

    def trace_any(self, rays: np.ndarray, counted: bool = False) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        occ = np.empty(rays.shape[0], np.uint8)
        fn = self.api._trace_any_counted if counted else self.api._trace_any
        self._ck(fn(self.h, rays.ctypes.data, rays.shape[0], occ.ctypes.data, 1, None), "trace_any")
        return occ

    # raw device-pointer variants (ints), for bench / torch tensors
    def trace_closest_dev(self, rays_ptr: int, n: int, hits_ptr: int, stream: int = 0):
        self._ck(self.api._trace_closest(self.h, rays_ptr, n, hits_ptr, 0, stream), "trace_closest")

    def trace_any_dev(self, rays_ptr: int, n: int, occ_ptr: int, stream: int = 0):
        self._ck(self.api._trace_any(self.h, rays_ptr, n, occ_ptr, 0, stream), "trace_any")

    # ---- render
    def render(self, width, height, max_depth, first_sample, sample_count, sample_stride=1, clear=True,
               want_aov=True):
        """Host-buffer render; returns (sum_rgba[h,w,4], albedo[h,w,3] | None, normal[h,w,3] | None)."""
        p = RenderParams(width, height, max_depth, first_sample, sample_count, sample_stride,
                         RENDER_HOST_BUFFERS | (RENDER_CLEAR if clear else 0), 0)
        acc = np.zeros((height, width, 4), np.float32)
        alb = np.zeros((height, width, 3), np.float32) if want_aov else None
        nrm = np.zeros((height, width, 3), np.float32) if want_aov else None
        self._ck(self.api._render(self.h, C.byref(p), acc.ctypes.data, alb.ctypes.data if want_aov else None,
                                  nrm.ctypes.data if want_aov else None, None), "render")
        return acc, alb, nrm

    def render_ex(self, width, height, max_depth, first_sample, sample_count, sample_stride=1, clear=True):
        """pb_render_ex with host buffers: returns (sum_rgba, albedo, normal, test[h,w])."""
        p = RenderParams(width, height, max_depth, first_sample, sample_count, sample_stride,
                         RENDER_HOST_BUFFERS | (RENDER_CLEAR if clear else 0), 0)
        acc = np.zeros((height, width, 4), np.float32)
        alb, nrm = np.zeros((height, width, 3), np.float32), np.zeros((height, width, 3), np.float32)
        tst = np.zeros((height, width), np.float32)
        t = RenderTargets(acc.ctypes.data, alb.ctypes.data, nrm.ctypes.data, tst.ctypes.data)
        self._ck(self.api._render_ex(self.h, C.byref(p), C.byref(t), None), "render_ex")
        return acc, alb, nrm, tst

    def render_dev(self, params: RenderParams, sum_ptr: int, albedo_ptr: int = 0, normal_ptr: int = 0, stream: int = 0):
        self._ck(self.api._render(self.h, C.byref(params), sum_ptr, albedo_ptr or None, normal_ptr or None, stream), "render")

    def render_into(self, params: RenderParams, acc: np.ndarray):
        """Host-buffer render accumulating into an existing host array."""
        self._ck(self.api._render(self.h, C.byref(params), acc.ctypes.data, None, None, None), "render")

    def tonemap(self, frame_rgba: np.ndarray, tone_mapping=True, gamma_correct=True) -> np.ndarray:
        """output.hlsl display transform of a resolved float4 frame (host arrays) -> uint8 RGBA."""
        a = np.ascontiguousarray(frame_rgba, np.float32)
        out = np.zeros(a.shape[:-1] + (4,), np.uint8)
        self._ck(self.api._tonemap(self.h, a.ctypes.data, out.ctypes.data, a.size // 4, int(tone_mapping), int(gamma_correct), 1, None), "tonemap")
        return out

    # ---- multi-GPU (one process per GPU): NCCL communicator of this context
    def comm_unique_id(self) -> bytes:
        buf = (C.c_ubyte * 128)()
        self._ck(self.api._comm_unique_id(buf), "comm_unique_id")
        return bytes(buf)

    def comm_init_rank(self, unique_id: bytes, n_ranks: int, rank: int):
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        self._ck(self.api._comm_init_rank(self.h, buf, n_ranks, rank), "comm_init_rank")

    def comm_info(self):
        n, r = C.c_int(), C.c_int()
        self._ck(self.api._comm_info(self.h, C.byref(n), C.byref(r)), "comm_info")
        return n.value, r.value

    def wait(self):
        """Synchronises renders enqueued with RENDER_ASYNC."""
        self._ck(self.api._wait(self.h), "wait")

    def stats(self) -> dict:
        s = Stats()
        self._ck(self.api._get_stats(self.h, C.byref(s)), "get_stats")
        return s.as_dict()


_product = None


def load() -> Api:
    """Load the CUDA product library.  Raises if it has not been built — never falls back."""
    global _product
    if _product is None:
        api = Api(LIB_PATH, "pb_")
        vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
        api._f("trace_closest_counted", i32, [vp, vp, u64, vp, i32, vp])
        api._f("trace_any_counted", i32, [vp, vp, u64, vp, i32, vp])
        api._f("export_bvh", i32, [vp, vp, C.POINTER(u64), vp, C.POINTER(u64)])
        api._f("trace_closest_static", i32, [vp, vp, u64, vp, i32, vp])
        api._f("comm_unique_id", i32, [vp])
        api._f("comm_init_rank", i32, [vp, vp, i32, i32])
        api._f("comm_info", i32, [vp, C.POINTER(i32), C.POINTER(i32)])
        api._f("comm_destroy", i32, [vp])
        api._f("create_multi", i32, [C.POINTER(vp), C.POINTER(i32), i32])
        api._f("render_multi", i32, [C.POINTER(vp), i32, C.POINTER(RenderParams), C.POINTER(RenderTargets), C.POINTER(vp)])
        _product = api
    return _product
