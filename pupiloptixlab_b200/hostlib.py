"""ctypes view of libpupil_host.so (include/pupil_host.h): the C++ host layer — mitsuba-style scene XML
loader, World::LoadScene hand-off and the headless System / B200PTPass loop.  Used by tests and tools;
the library itself is C++ (pupiloptixlab_b200/host/)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "host", "libpupil_host.so")

_vp, _u32, _i32 = C.c_void_p, C.c_uint32, C.c_int
_fpp = C.POINTER(C.c_float)


class Sink(C.Structure):
    _fields_ = [("scene_begin", C.CFUNCTYPE(_i32, _vp)),
                ("add_material", C.CFUNCTYPE(_i32, _vp, C.POINTER(capi.Material), C.POINTER(_u32))),
                ("add_trimesh", C.CFUNCTYPE(_i32, _vp, _fpp, _fpp, _fpp, _u32, C.POINTER(_u32), _u32, _fpp, _u32, _u32, C.POINTER(capi.Texture), C.POINTER(_u32))),
                ("add_sphere", C.CFUNCTYPE(_i32, _vp, _fpp, _u32, _u32, C.POINTER(capi.Texture), C.POINTER(_u32))),
                ("set_env", C.CFUNCTYPE(_i32, _vp, C.POINTER(capi.Env))),
                ("set_camera", C.CFUNCTYPE(_i32, _vp, _fpp, _fpp)),
                ("scene_commit", C.CFUNCTYPE(_i32, _vp))]


class SceneInfo(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("film_w", C.c_int32), ("film_h", C.c_int32), ("fov_y", C.c_float),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("n_shapes", _u32), ("n_emissive_shapes", _u32),
                ("n_triangles", _u32), ("n_spheres", _u32), ("has_env", _u32), ("n_warnings", _u32)]


class HostError(RuntimeError):
    pass


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HostError(f"{LIB_PATH} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        lib = C.CDLL(LIB_PATH)
        lib.ph_last_error.restype = C.c_char_p
        lib.ph_scene_warning.restype = C.c_char_p
        lib.ph_scene_warning.argtypes = [_vp, _u32]
        lib.ph_scene_load_xml.argtypes = [C.c_char_p, C.POINTER(_vp)]
        lib.ph_scene_load_xml_string.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(_vp)]
        lib.ph_scene_free.argtypes = [_vp]
        lib.ph_scene_info_get.argtypes = [_vp, C.POINTER(SceneInfo)]
        lib.ph_scene_camera.argtypes = [_vp, _fpp, _fpp]
        lib.ph_scene_shape.argtypes = [_vp, _u32, C.POINTER(_u32), _fpp, C.POINTER(capi.Material), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(capi.Texture)]
        lib.ph_scene_upload.argtypes = [_vp, C.POINTER(Sink), _vp]
        lib.ph_default_sink.restype = C.POINTER(Sink)
        lib.ph_run_path_tracer.argtypes = [C.c_char_p, _i32, _u32, _vp, _vp, _vp, C.POINTER(_u32), C.POINTER(_u32)]
        _lib = lib
    return _lib


def sink_for(api: capi.Api) -> Sink:
    """A ph_sink made of the scene-ingestion entry points of `api` (pb_* of the product, orc_* of the test oracle)."""
    s = Sink()
    for name, _ in Sink._fields_:
        fn = getattr(api.lib, api.prefix + name)
        setattr(s, name, C.cast(fn, dict(Sink._fields_)[name]))
    return s


class Scene:
    """resource::Scene loaded from XML (file path or text)."""

    def __init__(self, path: str | None = None, text: str | None = None, root_dir: str = ""):
        self.lib = load()
        self.h = _vp()
        if path is not None:
            rc = self.lib.ph_scene_load_xml(path.encode(), C.byref(self.h))
        else:
            rc = self.lib.ph_scene_load_xml_string(text.encode(), root_dir.encode(), C.byref(self.h))
        if rc != 0:
            raise HostError(self.lib.ph_last_error().decode())

    def __del__(self):
        try:
            if self.h:
                self.lib.ph_scene_free(self.h)
                self.h = _vp()
        except Exception:
            pass

    def info(self) -> SceneInfo:
        i = SceneInfo()
        self.lib.ph_scene_info_get(self.h, C.byref(i))
        return i

    def warnings(self) -> list[str]:
        return [self.lib.ph_scene_warning(self.h, k).decode() for k in range(self.info().n_warnings)]

    def camera(self):
        a, b = np.zeros(16, np.float32), np.zeros(16, np.float32)
        self.lib.ph_scene_camera(self.h, capi._fp(a), capi._fp(b))
        return a.reshape(4, 4), b.reshape(4, 4)

    def shape(self, i: int) -> dict:
        ty, fl, em = _u32(), _u32(), _u32()
        tw = np.zeros(12, np.float32)
        m, rad = capi.Material(), capi.Texture()
        if self.lib.ph_scene_shape(self.h, i, C.byref(ty), capi._fp(tw), C.byref(m), C.byref(fl), C.byref(em), C.byref(rad)) != 0:
            raise IndexError(i)
        return dict(type=("obj", "sphere", "cube", "rectangle")[ty.value], to_world=tw.reshape(3, 4), material=m, flags=fl.value,
                    is_emitter=bool(em.value), radiance=rad)

    def upload(self, ctx: capi.Context):
        """World::LoadScene into `ctx` (a capi.Context of the product or of the test oracle)."""
        sink = sink_for(ctx.api)
        rc = self.lib.ph_scene_upload(self.h, C.byref(sink), ctx.h)
        if rc != 0:
            raise HostError(f"{self.lib.ph_last_error().decode()}: {ctx.api._last_error(ctx.h).decode()}")


class Session:
    """ph_session: a headless System + B200PTPass kept alive between calls (render, move an object, render, read)."""
    BUFFERS = {"final result": 4, "pt accum buffer": 4, "albedo": 3, "normal": 3, "test": 1}

    def __init__(self, xml_path: str, device: int = 0):
        self.lib = load()
        L = self.lib
        L.ph_session_create.argtypes = [C.c_char_p, _i32, C.POINTER(_vp)]
        L.ph_session_destroy.argtypes = [_vp]
        L.ph_session_run.argtypes = [_vp, _u32]
        L.ph_session_move_object.argtypes = [_vp, _u32, _fpp]
        L.ph_session_read.argtypes = [_vp, C.c_char_p, _vp, C.c_uint64]
        L.ph_session_film.argtypes = [_vp, C.POINTER(_u32), C.POINTER(_u32)]
        L.ph_session_sample_count.argtypes = [_vp]
        L.ph_session_sample_count.restype = _u32
        self.h = _vp()
        rc = L.ph_session_create(xml_path.encode(), device, C.byref(self.h))
        if rc != 0:
            raise HostError(f"ph_session_create failed ({rc}): {L.ph_last_error().decode()}")
        w, h = _u32(), _u32()
        L.ph_session_film(self.h, C.byref(w), C.byref(h))
        self.w, self.h_px = w.value, h.value

    def close(self):
        if self.h:
            self.lib.ph_session_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run(self, frames: int):
        rc = self.lib.ph_session_run(self.h, frames)
        if rc != 0:
            raise HostError(f"ph_session_run failed ({rc}): {self.lib.ph_last_error().decode()}")

    def move_object(self, index: int, to_world):
        m = np.ascontiguousarray(to_world, np.float32).reshape(16).copy()
        rc = self.lib.ph_session_move_object(self.h, index, capi._fp(m))
        if rc != 0:
            raise HostError(f"ph_session_move_object failed ({rc}): {self.lib.ph_last_error().decode()}")

    def read(self, name: str) -> np.ndarray:
        ch = self.BUFFERS[name]
        out = np.zeros((self.h_px, self.w, ch) if ch > 1 else (self.h_px, self.w), np.float32)
        rc = self.lib.ph_session_read(self.h, name.encode(), out.ctypes.data, out.nbytes)
        if rc != 0:
            raise HostError(f"ph_session_read({name}) failed ({rc}): {self.lib.ph_last_error().decode()}")
        return out

    def sample_count(self) -> int:
        return int(self.lib.ph_session_sample_count(self.h))


def load_image(path: str) -> np.ndarray:
    """BitmapTexture::Load: (h, w, 4) float32 texels, row 0 = top."""
    lib = load()
    p, w, h = C.POINTER(C.c_float)(), _u32(), _u32()
    lib.ph_image_load.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.ph_image_free.argtypes = [C.POINTER(C.c_float)]
    if lib.ph_image_load(path.encode(), C.byref(p), C.byref(w), C.byref(h)) != 0:
        raise HostError(lib.ph_last_error().decode())
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    lib.ph_image_free(p)
    return out


def save_image(path: str, rgba: np.ndarray):
    """BitmapTexture::Save by extension (.hdr / .exr as the reference writes them, .pfm); uint8 input -> 8-bit PNG."""
    lib = load()
    a = np.ascontiguousarray(rgba)
    h, w = a.shape[:2]
    if a.dtype == np.uint8:
        lib.ph_image_save_png8.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        rc = lib.ph_image_save_png8(path.encode(), a.ctypes.data, w, h)
    else:
        a = np.ascontiguousarray(a, np.float32)
        lib.ph_image_save.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        rc = lib.ph_image_save(path.encode(), a.ctypes.data, w, h)
    if rc != 0:
        raise HostError(lib.ph_last_error().decode())


def run_path_tracer(xml_path: str, frames: int, device: int = 0, width: int | None = None, height: int | None = None):
    """Headless System + B200PTPass: returns (final_rgba[h,w,4] running mean, albedo[h,w,3], normal[h,w,3])."""
    lib = load()
    info = Scene(path=xml_path).info()
    w, h = info.film_w, info.film_h
    fin, alb, nrm = np.zeros((h, w, 4), np.float32), np.zeros((h, w, 3), np.float32), np.zeros((h, w, 3), np.float32)
    ow, oh = _u32(), _u32()
    rc = lib.ph_run_path_tracer(xml_path.encode(), device, frames, fin.ctypes.data, alb.ctypes.data, nrm.ctypes.data, C.byref(ow), C.byref(oh))
    if rc != 0:
        raise HostError(f"ph_run_path_tracer failed ({rc}): {lib.ph_last_error().decode()}")
    assert (ow.value, oh.value) == (w, h)
    return fin, alb, nrm
