"""Procedural inputs for tests and benchmarks: camera matrices, the reference's built-in meshes,
the Cornell-box configs (C1/C2), the displaced-mesh + spheres scene (C3/C4/C5) and ray sets.

Everything here only *describes* a scene through the C ABI (`capi.Context`); it works unchanged on
the CUDA product and on the test oracle because both export the same entry points.
"""
from __future__ import annotations

import numpy as np

from . import capi

# ------------------------------------------------------------------------------------------------
# camera / transform math (restated from the public DirectXMath definitions)


def perspective_sample_to_camera(fov_y_deg: float, aspect: float, near: float = 0.01, far: float = 10000.0) -> np.ndarray:
    """util::Camera::GetSampleToCameraMatrix (framework/util/camera.cpp:7-20):
    transpose(inverse(PerspectiveFovRH * Translation(1,1,0) * Scaling(.5,.5,1))), row-vector maths."""
    fov = np.float32(fov_y_deg) / np.float32(180.0) * np.float32(3.14159265358979323846)
    h = 1.0 / np.tan(np.float64(fov) * 0.5)
    w = h / np.float64(aspect)
    r = far / (near - far)
    proj = np.array([[w, 0, 0, 0], [0, h, 0, 0], [0, 0, r, -1], [0, 0, r * near, 0]], np.float64)
    T = np.eye(4)
    T[3, :3] = (1, 1, 0)
    S = np.diag([0.5, 0.5, 1.0, 1.0])
    M = proj @ T @ S
    return np.linalg.inv(M).T.astype(np.float32)


def fov_x_to_y(fov_deg: float, width: int, height: int) -> float:
    """framework/resource/scene.cpp:122-127 (fov_axis = x)."""
    aspect = np.float32(height) / np.float32(width)
    radian = np.float32(fov_deg) * np.float32(3.14159265358979323846) / np.float32(180.0) * np.float32(0.5)
    t = np.float32(np.tan(radian)) * aspect
    return float(np.float32(2.0) * np.float32(np.arctan(t)) * np.float32(180.0) / np.float32(3.14159265358979323846))


def _flip_xz(m: np.ndarray) -> np.ndarray:
    m = m.copy()
    m[:3, 0] *= -1
    m[:3, 2] *= -1
    return m


def lookat_to_world(origin, target, up) -> np.ndarray:
    """Transform::LookAt (framework/util/transform.cpp:86-98) followed by the column flip of
    LoadTransform3D (framework/resource/xml/util_loader.cpp:161-166)."""
    e, f, u = (np.asarray(v, np.float64) for v in (origin, target, up))
    z = e - f
    z /= np.linalg.norm(z)
    x = np.cross(u, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = x, y, z, e
    return _flip_xz(m).astype(np.float32)


def sensor_to_world(to_world: np.ndarray) -> np.ndarray:
    """The sensor case flips columns 0 and 2 (again) — framework/resource/scene.cpp:134-139."""
    return _flip_xz(np.asarray(to_world, np.float32).reshape(4, 4))


def set_perspective_camera(ctx: capi.Context, fov_deg, width, height, to_world_xml, fov_axis="x"):
    fov_y = fov_x_to_y(fov_deg, width, height) if fov_axis == "x" else fov_deg
    s2c = perspective_sample_to_camera(fov_y, np.float32(width) / np.float32(height))
    c2w = sensor_to_world(to_world_xml)
    ctx.set_camera(s2c, c2w)
    return s2c, c2w


def scale(sx, sy, sz):
    return np.diag([sx, sy, sz, 1.0]).astype(np.float32)


def translate(x, y, z):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = (x, y, z)
    return m


def rotate(ux, uy, uz, angle_deg):
    """Transform::Rotate quaternion form (framework/util/transform.cpp:7-47)."""
    n = np.sqrt(ux * ux + uy * uy + uz * uz)
    ux, uy, uz = ux / n, uy / n, uz / n
    th = angle_deg / 180.0 * np.pi
    a, s = np.cos(0.5 * th), np.sin(0.5 * th)
    b, c, d = s * ux, s * uy, s * uz
    m = np.eye(4)
    m[:3, :3] = [[1 - 2 * c * c - 2 * d * d, 2 * b * c - 2 * a * d, 2 * a * c + 2 * b * d],
                 [2 * b * c + 2 * a * d, 1 - 2 * b * b - 2 * d * d, 2 * c * d - 2 * a * b],
                 [2 * b * d - 2 * a * c, 2 * a * b + 2 * c * d, 1 - 2 * b * b - 2 * c * c]]
    return m.astype(np.float32)


# ------------------------------------------------------------------------------------------------
# built-in meshes (framework/resource/shape.cpp:20-66)

RECT_POS = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], np.float32)
RECT_NRM = np.tile(np.array([[0, 0, 1]], np.float32), (4, 1))
RECT_UV = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
RECT_IDX = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)


def _cube():
    faces = [  # (normal, 4 corners) in the reference's vertex order
        ((-1, 0, 0), [(-1, -1, -1), (-1, -1, 1), (-1, 1, 1), (-1, 1, -1)]),
        ((0, 0, -1), [(1, -1, -1), (-1, -1, -1), (-1, 1, -1), (1, 1, -1)]),
        ((1, 0, 0), [(1, -1, 1), (1, -1, -1), (1, 1, -1), (1, 1, 1)]),
        ((0, 0, 1), [(-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)]),
        ((0, 1, 0), [(-1, 1, 1), (1, 1, 1), (1, 1, -1), (-1, 1, -1)]),
        ((0, -1, 0), [(-1, -1, -1), (1, -1, -1), (1, -1, 1), (-1, -1, 1)]),
    ]
    pos = np.array([c for _, cs in faces for c in cs], np.float32)
    nrm = np.array([n for n, cs in faces for _ in cs], np.float32)
    uv = np.tile(RECT_UV, (6, 1))
    idx = np.array([[4 * f + a, 4 * f + b, 4 * f + c] for f in range(6) for a, b, c in ((0, 1, 2), (0, 2, 3))], np.uint32)
    return pos, nrm, uv, idx


CUBE_POS, CUBE_NRM, CUBE_UV, CUBE_IDX = _cube()


def add_rectangle(ctx, to_world, material, flags=0, area_radiance=None):
    return ctx.add_trimesh(RECT_POS, RECT_NRM, RECT_UV, RECT_IDX, to_world, flags, material, area_radiance)


def add_cube(ctx, to_world, material, flags=0, area_radiance=None):
    return ctx.add_trimesh(CUBE_POS, CUBE_NRM, CUBE_UV, CUBE_IDX, to_world, flags, material, area_radiance)


def add_sphere(ctx, center, radius, material, to_world=None, flags=0, area_radiance=None):
    """ShapeLoader<_sphere> (shape.cpp:106-125,196-197): to_world * Translate(center) * Scale(radius)."""
    m = translate(*center) @ scale(radius, radius, radius)
    if to_world is not None:
        m = np.asarray(to_world, np.float32).reshape(4, 4) @ m
    return ctx.add_sphere(m.astype(np.float32), flags, material, area_radiance)


# ------------------------------------------------------------------------------------------------
# Cornell box: geometry of data/static/cornellbox.xml (the matrices are the XML's own values)

def _m(s):
    return np.array([float(x) for x in s.split()], np.float32).reshape(4, 4)


CORNELL_SENSOR = _m("-1 0 0 0 0 1 0 1 0 0 -1 6.8 0 0 0 1")
CORNELL = {
    "Floor": ("rectangle", _m("-4.37114e-008 1 4.37114e-008 0 0 -8.74228e-008 2 0 1 4.37114e-008 1.91069e-015 0 0 0 0 1")),
    "Ceiling": ("rectangle", _m("-1 7.64274e-015 -1.74846e-007 0 8.74228e-008 8.74228e-008 -2 2 0 -1 -4.37114e-008 0 0 0 0 1")),
    "BackWall": ("rectangle", _m("1.91069e-015 1 1.31134e-007 0 1 3.82137e-015 -8.74228e-008 1 -4.37114e-008 1.31134e-007 -2 -1 0 0 0 1")),
    "RightWall": ("rectangle", _m("4.37114e-008 -1.74846e-007 2 1 1 3.82137e-015 -8.74228e-008 1 3.82137e-015 1 2.18557e-007 0 0 0 0 1")),
    "LeftWall": ("rectangle", _m("-4.37114e-008 8.74228e-008 -2 -1 1 3.82137e-015 -8.74228e-008 1 0 -1 -4.37114e-008 0 0 0 0 1")),
    "ShortBox": ("cube", _m("0.0851643 0.289542 1.31134e-008 0.328631 3.72265e-009 1.26563e-008 -0.3 0.3 -0.284951 0.0865363 5.73206e-016 0.374592 0 0 0 1")),
    "TallBox": ("cube", _m("0.286776 0.098229 -2.29282e-015 -0.335439 -4.36233e-009 1.23382e-008 -0.6 0.6 -0.0997984 0.282266 2.62268e-008 -0.291415 0 0 0 1")),
    "Light": ("rectangle", _m("0.235 -1.66103e-008 -7.80685e-009 -0.005 -2.05444e-008 3.90343e-009 -0.0893 1.98 2.05444e-008 0.19 8.30516e-009 -0.03 0 0 0 1")),
}
WHITE = (0.725, 0.71, 0.68)


def cornell_box(ctx: capi.Context, width=256, height=256, variant="diffuse") -> dict:
    """C1 (variant='diffuse'): cornellbox.xml as shipped.  C2 (variant='materials'): walls keep their
    diffuse BSDFs, the boxes and five added spheres carry the other six materials with the
    parameters of data/static/material_test.xml:30-114."""
    ctx.scene_begin()
    M = capi.make_material
    diff = lambda rgb: ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True, reflectance=rgb))
    mats = {"Floor": diff(WHITE), "Ceiling": diff(WHITE), "BackWall": diff(WHITE),
            "RightWall": diff((0.14, 0.45, 0.091)), "LeftWall": diff((0.63, 0.065, 0.05)),
            "ShortBox": diff(WHITE), "TallBox": diff(WHITE), "Light": diff((0, 0, 0))}
    if variant == "materials":
        mats["ShortBox"] = ctx.add_material(M(capi.MAT_ROUGH_PLASTIC, int_ior=1.5, ext_ior=1.0, alpha=0.35,
                                              reflectance=(0.647814,) * 3))
        mats["TallBox"] = ctx.add_material(M(capi.MAT_ROUGH_CONDUCTOR, alpha=0.35, specular_reflectance=0.3,
                                             eta=(0.200438, 0.924033, 1.10221), k=(3.91295, 2.45285, 2.14219)))
        mats["Floor"] = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True,
                                           reflectance=capi.checker_texture((0.725, 0.71, 0.68), (0.3, 0.3, 0.32), 8, 8)))
    textured = variant in ("textured", "envmap")
    if textured:
        # bitmap textures in every role the reference samples them (cuda/texture.h:52-54): reflectance with linear /
        # wrap and point / mirror sampling, a roughness map, a bitmap-radiance area light (emitter.cpp:85-100)
        mats["BackWall"] = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True,
                                              reflectance=capi.bitmap_texture(procedural_bitmap(48, 32, 1), capi.ADDRESS_WRAP, capi.FILTER_LINEAR, 2, 2)))
        mats["Floor"] = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True,
                                           reflectance=capi.bitmap_texture(procedural_bitmap(16, 16, 2), capi.ADDRESS_MIRROR, capi.FILTER_POINT, 3, 3)))
        mats["TallBox"] = ctx.add_material(M(capi.MAT_ROUGH_CONDUCTOR, specular_reflectance=0.8, eta=(0.200438, 0.924033, 1.10221),
                                             k=(3.91295, 2.45285, 2.14219),
                                             alpha=capi.bitmap_texture(0.05 + 0.5 * procedural_bitmap(8, 8, 3), capi.ADDRESS_CLAMP, capi.FILTER_LINEAR)))
        mats["ShortBox"] = ctx.add_material(M(capi.MAT_ROUGH_PLASTIC, int_ior=1.5, ext_ior=1.0, alpha=0.2,
                                              reflectance=capi.bitmap_texture(procedural_bitmap(32, 32, 4), capi.ADDRESS_WRAP, capi.FILTER_LINEAR)))
    objs = {}
    for name, (kind, m) in CORNELL.items():
        if variant == "envmap" and name in ("Ceiling", "Light"):
            continue      # open top: the box is lit by the environment map only
        ar = capi.rgb_texture(17, 12, 4) if name == "Light" else None
        if textured and name == "Light":
            ar = capi.bitmap_texture(procedural_bitmap(8, 8, 5) * np.float32(30.0), capi.ADDRESS_CLAMP, capi.FILTER_LINEAR)
        add = add_rectangle if kind == "rectangle" else add_cube
        objs[name] = add(ctx, m, mats[name], 0, ar)
    if variant == "envmap":
        rot = np.array([[0.8, 0, 0.6, 0], [0, 1, 0, 0], [-0.6, 0, 0.8, 0]], np.float32)   # y-rotation: exercises to_world / to_local
        ctx.set_env(capi.env_map(procedural_envmap(64, 32), scale=1.5, to_world=rot))
    if variant == "materials":
        sph = [
            ((-0.55, 0.25, 0.55), 0.25, M(capi.MAT_DIELECTRIC, int_ior=1.5, ext_ior=1.0)),
            ((0.0, 0.2, 0.7), 0.2, M(capi.MAT_ROUGH_DIELECTRIC, int_ior=1.5, ext_ior=1.0, alpha=0.35)),
            ((0.6, 0.2, 0.65), 0.2, M(capi.MAT_PLASTIC, int_ior=1.5, ext_ior=1.0, reflectance=(0.647814,) * 3)),
            ((0.33, 0.78, 0.37), 0.18, M(capi.MAT_CONDUCTOR, specular_reflectance=0.3,
                                         eta=(0.200438, 0.924033, 1.10221), k=(3.91295, 2.45285, 2.14219))),
            ((-0.34, 1.38, -0.29), 0.18, M(capi.MAT_ROUGH_CONDUCTOR, alpha=0.95, specular_reflectance=0.8,
                                           eta=(0.15494, 0.11648, 0.13809), k=(4.81810, 3.11562, 2.14240))),
        ]
        for i, (c, r, mat) in enumerate(sph):
            objs[f"sphere{i}"] = add_sphere(ctx, c, r, ctx.add_material(mat))
        # a small emissive sphere exercises SphereEmitter (emitter/sphere.h)
        objs["bulb"] = add_sphere(ctx, (0.7, 1.6, 0.4), 0.06, mats["Light"], area_radiance=capi.rgb_texture(30, 30, 30))
    set_perspective_camera(ctx, 19.5, width, height, CORNELL_SENSOR)
    ctx.commit()
    return {"width": width, "height": height, "objects": objs, "materials": mats}


def edge_scene(ctx: capi.Context, width=160, height=120, env=True, emitters=True) -> dict:
    """Branches of the hot path that the Cornell variants do not reach (parity gaps):
    constant environment emitter (emitter/env.h:67-86), flip_normals / flip_tex_coords (geometry.h:62-75; the latter is
    the default of every imported .obj), meshes without vertex normals (cross-product branch, geometry.h:57-60) and
    without uvs, nonlinear plastic / rough plastic (plastic.h:38, rough_plastic.h:44), a dielectric with
    int_ior < ext_ior (eta < 1), non-uniformly scaled spheres (one of them an area emitter, sphere.h:14-42).
    `emitters=False` leaves the scene without any light (NEE has nothing to select)."""
    ctx.scene_begin()
    M = capi.make_material
    chk = capi.checker_texture((0.8, 0.25, 0.1), (0.15, 0.3, 0.85), 6, 6)
    m_floor = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True, reflectance=chk))
    m_wall = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True, reflectance=(0.7, 0.68, 0.6)))
    m_one_sided = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=False, reflectance=(0.3, 0.6, 0.35)))
    m_plastic_nl = ctx.add_material(M(capi.MAT_PLASTIC, twosided=True, nonlinear=True, int_ior=1.5, ext_ior=1.0, reflectance=chk,
                                      specular_reflectance=(0.9, 0.9, 0.8)))
    m_rplastic_nl = ctx.add_material(M(capi.MAT_ROUGH_PLASTIC, twosided=True, nonlinear=True, alpha=0.15, int_ior=1.49, ext_ior=1.0,
                                       reflectance=(0.5, 0.5, 0.1)))
    m_glass_inv = ctx.add_material(M(capi.MAT_DIELECTRIC, int_ior=1.0, ext_ior=1.5))
    m_rglass_inv = ctx.add_material(M(capi.MAT_ROUGH_DIELECTRIC, int_ior=1.0, ext_ior=1.33, alpha=0.2))
    m_black = ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True, reflectance=(0, 0, 0)))
    objs = {}
    # floor with uvs, v flipped (ShapeInstance::flip_tex_coords)
    objs["floor"] = ctx.add_trimesh(RECT_POS, RECT_NRM, RECT_UV, RECT_IDX, rotate(1, 0, 0, -90) @ scale(1.5, 1.5, 1), capi.SHAPE_FLIP_TEX_COORDS, m_floor)
    # back wall without normals and without uvs
    objs["back"] = ctx.add_trimesh(RECT_POS, None, None, RECT_IDX, translate(0, 1, -1.5) @ scale(1.5, 1, 1), 0, m_wall)
    # left wall: one-sided material, geometric normal points away from the room, flip_normals turns it back
    objs["left"] = ctx.add_trimesh(RECT_POS, RECT_NRM, RECT_UV, RECT_IDX, translate(-1.5, 1, 0) @ rotate(0, 1, 0, -90) @ scale(1.5, 1, 1),
                                   capi.SHAPE_FLIP_NORMALS, m_one_sided)
    # cube without normals (uvs kept): nonlinear plastic with a checkerboard reflectance
    objs["cube"] = ctx.add_trimesh(CUBE_POS, None, CUBE_UV, CUBE_IDX, translate(-0.7, 0.3, -0.4) @ rotate(0, 1, 0, 30) @ scale(0.3, 0.3, 0.3),
                                   capi.SHAPE_FLIP_TEX_COORDS, m_plastic_nl)
    # a cube with normals but no uvs, flip_normals + two-sided material
    objs["cube2"] = ctx.add_trimesh(CUBE_POS, CUBE_NRM, None, CUBE_IDX, translate(0.9, 0.2, -0.7) @ rotate(0, 1, 0, -20) @ scale(0.2, 0.2, 0.2),
                                    capi.SHAPE_FLIP_NORMALS, m_rplastic_nl)
    objs["glass"] = add_sphere(ctx, (0.1, 0.35, 0.2), 0.35, m_glass_inv)
    # non-uniformly scaled, rotated spheres
    objs["ellipsoid"] = ctx.add_sphere(translate(0.85, 0.22, 0.45) @ rotate(1, 1, 0, 40) @ scale(0.35, 0.15, 0.22), 0, m_rplastic_nl)
    objs["ellipsoid_glass"] = ctx.add_sphere(translate(-0.55, 0.9, 0.1) @ rotate(0, 0, 1, 25) @ scale(0.3, 0.12, 0.2), capi.SHAPE_FLIP_NORMALS, m_rglass_inv)
    if emitters:
        objs["ellipsoid_light"] = ctx.add_sphere(translate(-0.2, 1.6, -0.6) @ scale(0.12, 0.05, 0.09), 0, m_black, capi.rgb_texture(25, 22, 18))
        objs["light"] = ctx.add_trimesh(RECT_POS, RECT_NRM, RECT_UV, RECT_IDX, translate(0.6, 1.9, 0.0) @ rotate(1, 0, 0, 90) @ scale(0.2, 0.2, 1), 0, m_black,
                                        capi.rgb_texture(12, 12, 14))
        if env:
            ctx.set_env(capi.env_const(0.45, 0.55, 0.8))
    to_world = lookat_to_world((0.2, 1.3, 3.4), (0.0, 0.6, 0.0), (0, 1, 0))
    set_perspective_camera(ctx, 42.0, width, height, to_world)
    ctx.commit()
    return {"width": width, "height": height, "objects": objs}


def procedural_bitmap(w: int, h: int, seed: int) -> np.ndarray:
    """(h, w, 4) float RGBA test pattern in [0.05, 0.95]: smooth gradients + seeded blocks (no image files)."""
    rng = np.random.RandomState(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.empty((h, w, 4), np.float32)
    img[..., 0] = 0.5 + 0.45 * np.sin(2 * np.pi * (x / w + 0.13 * seed))
    img[..., 1] = 0.5 + 0.45 * np.cos(2 * np.pi * (y / h) * 2 + seed)
    blocks = rng.uniform(0.05, 0.95, ((h + 3) // 4, (w + 3) // 4)).astype(np.float32)
    img[..., 2] = np.kron(blocks, np.ones((4, 4), np.float32))[:h, :w]
    img[..., 3] = 1.0
    return np.clip(img, 0.05, 0.95).astype(np.float32)


def procedural_envmap(w: int, h: int) -> np.ndarray:
    """(h, w, 4) lat-long sky: blue gradient, warm horizon band and a bright 'sun' patch (so that the row / column
    CDFs of BuildEnvMapCdfTable, world/emitter.cpp:107-149, are far from uniform)."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    theta = (y + 0.5) / h
    sky = np.stack([0.25 + 0.3 * theta, 0.35 + 0.35 * theta, 0.9 - 0.3 * theta], -1)
    sky[np.abs(theta - 0.5) < 0.06] += np.array([0.6, 0.35, 0.1], np.float32)
    sun = ((x - 0.3 * w) ** 2 / (0.04 * w) ** 2 + (y - 0.22 * h) ** 2 / (0.07 * h) ** 2) < 1.0
    sky[sun] = (40.0, 36.0, 28.0)
    sky[theta > 0.55] *= 0.15         # dark ground half
    return np.concatenate([sky, np.ones((h, w, 1))], -1).astype(np.float32)


# ------------------------------------------------------------------------------------------------
# Procedural displaced mesh + spheres (configs C3/C4/C5)

def _hash2(ix, iy, seed):
    h = (ix.astype(np.uint64) * np.uint64(0x9E3779B1) + iy.astype(np.uint64) * np.uint64(0x85EBCA77) + np.uint64(seed) * np.uint64(0xC2B2AE3D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    return (h & np.uint64(0xFFFFFF)).astype(np.float64) / float(1 << 24)


def _value_noise(x, y, seed):
    x0, y0 = np.floor(x), np.floor(y)
    fx, fy = x - x0, y - y0
    ix, iy = x0.astype(np.int64) & 0xFFFF, y0.astype(np.int64) & 0xFFFF
    sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
    a, b = _hash2(ix, iy, seed), _hash2(ix + 1, iy, seed)
    c, d = _hash2(ix, iy + 1, seed), _hash2(ix + 1, iy + 1, seed)
    return (a + (b - a) * sx) * (1 - sy) + (c + (d - c) * sx) * sy


def height_field(x, z, seed=1234):
    """4 octaves of value noise, amplitude 0.15 * 2^-o."""
    h = np.zeros_like(x, dtype=np.float64)
    for o in range(4):
        f = 3.0 * (2 ** o)
        h += 0.15 * 2.0 ** (-o) * _value_noise((x + 1.0) * f, (z + 1.0) * f, seed + o)
    return h


def displaced_mesh(n: int, seed: int = 1234):
    """(n+1)^2 grid over [-1,1]^2 in XZ, 2 n^2 triangles; vertex normals from central differences,
    uv = grid coordinates.  n = 707 -> 1.0 M, 2236 -> 10.0 M, 7071 -> 100.0 M triangles.
    PB_MESH_CACHE=<dir> keeps the arrays on disk between runs (bench convenience only)."""
    import os
    cache = os.environ.get("PB_MESH_CACHE")
    path = os.path.join(cache, f"mesh_{n}_{seed}.npz") if cache else None
    if path and os.path.exists(path):
        z = np.load(path)
        return z["pos"], z["nrm"], z["uv"], z["idx"]
    mesh = _displaced_mesh(n, seed)
    if path:
        os.makedirs(cache, exist_ok=True)
        np.savez(path, pos=mesh[0], nrm=mesh[1], uv=mesh[2], idx=mesh[3])
    return mesh


def _displaced_mesh(n: int, seed: int = 1234):
    g = np.linspace(-1.0, 1.0, n + 1)
    X, Z = np.meshgrid(g, g, indexing="xy")
    Y = height_field(X, Z, seed)
    pos = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    e = 2.0 / n
    dx = (height_field(X + e, Z, seed) - height_field(X - e, Z, seed)) / (2 * e)
    dz = (height_field(X, Z + e, seed) - height_field(X, Z - e, seed)) / (2 * e)
    nrm = np.stack([-dx, np.ones_like(dx), -dz], -1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    nrm = nrm.reshape(-1, 3).astype(np.float32)
    uv = np.stack([(X + 1) * 0.5, (Z + 1) * 0.5], -1).reshape(-1, 2).astype(np.float32)
    i, j = np.meshgrid(np.arange(n, dtype=np.uint32), np.arange(n, dtype=np.uint32), indexing="xy")
    v00 = (j * (n + 1) + i).reshape(-1)
    v10, v01, v11 = v00 + 1, v00 + (n + 1), v00 + (n + 2)
    idx = np.empty((2 * n * n, 3), np.uint32)
    idx[0::2] = np.stack([v00, v01, v11], -1)   # counter-clockwise seen from +Y
    idx[1::2] = np.stack([v00, v11, v10], -1)
    return pos, nrm, uv, idx


MESH_N = {"1M": 707, "10M": 2236, "100M": 7071}
BENCH_CAMERA = dict(origin=(0.0, 1.2, 2.6), target=(0.0, 0.1, 0.0), up=(0.0, 1.0, 0.0), fov=45.0)


def seven_materials(ctx):
    """One material of each type (parameters as in data/static/material_test.xml)."""
    M = capi.make_material
    au = dict(eta=(0.200438, 0.924033, 1.10221), k=(3.91295, 2.45285, 2.14219))
    return [
        ctx.add_material(M(capi.MAT_DIFFUSE, twosided=True, reflectance=(0.7, 0.7, 0.7))),
        ctx.add_material(M(capi.MAT_DIELECTRIC, int_ior=1.5, ext_ior=1.0)),
        ctx.add_material(M(capi.MAT_ROUGH_DIELECTRIC, int_ior=1.5, ext_ior=1.0, alpha=0.35)),
        ctx.add_material(M(capi.MAT_CONDUCTOR, twosided=True, specular_reflectance=0.8, **au)),
        ctx.add_material(M(capi.MAT_ROUGH_CONDUCTOR, twosided=True, alpha=0.35, specular_reflectance=0.8, **au)),
        ctx.add_material(M(capi.MAT_PLASTIC, twosided=True, int_ior=1.5, ext_ior=1.0, reflectance=(0.65, 0.3, 0.2))),
        ctx.add_material(M(capi.MAT_ROUGH_PLASTIC, twosided=True, int_ior=1.5, ext_ior=1.0, alpha=0.35, reflectance=(0.2, 0.4, 0.65))),
    ]


def bench_scene(ctx, mesh_n: int, n_spheres: int = 10000, width=1920, height=1080, patchwork=True, mesh=None):
    """Displaced mesh (+ `n_spheres` spheres cycling the 7 materials) lit by one rectangle light at
    y = 1.5 (radiance 17,12,4; 0.5 x 0.5).  With `patchwork` the mesh is split into 7 objects by
    64x64-quad tiles (material = hash(tile) mod 7), otherwise it is one diffuse object (C3)."""
    pos, nrm, uv, idx = mesh if mesh is not None else displaced_mesh(mesh_n)
    ctx.scene_begin()
    mats = seven_materials(ctx)
    ident = np.eye(4, dtype=np.float32)
    if patchwork:
        quad = np.arange(mesh_n * mesh_n, dtype=np.int64)
        ti, tj = (quad % mesh_n) // 64, (quad // mesh_n) // 64
        m_of_quad = ((ti * 73856093) ^ (tj * 19349663)) % 7
        m_of_tri = np.repeat(m_of_quad, 2)
        for k in range(7):
            sel = idx[m_of_tri == k]
            if len(sel):
                ctx.add_trimesh(pos, nrm, uv, sel, ident, 0, mats[k])
    else:
        ctx.add_trimesh(pos, nrm, uv, idx, ident, 0, mats[0])
    rng = np.random.RandomState(42)
    ctr = np.stack([rng.uniform(-1, 1, n_spheres), rng.uniform(0.05, 0.6, n_spheres), rng.uniform(-1, 1, n_spheres)], -1)
    rad = rng.uniform(0.005, 0.02, n_spheres)
    for s in range(n_spheres):
        add_sphere(ctx, ctr[s], float(rad[s]), mats[s % 7])
    light_black = ctx.add_material(capi.make_material(capi.MAT_DIFFUSE, twosided=True, reflectance=(0, 0, 0)))
    lm = translate(0, 1.5, 0) @ rotate(1, 0, 0, 90) @ scale(0.25, 0.25, 1)
    add_rectangle(ctx, lm, light_black, 0, capi.rgb_texture(17, 12, 4))
    cam = BENCH_CAMERA
    to_world = lookat_to_world(cam["origin"], cam["target"], cam["up"])  # <lookat> flip; the sensor flips it again
    set_perspective_camera(ctx, cam["fov"], width, height, to_world)
    ctx.commit()
    return dict(width=width, height=height, n_tris=len(idx), n_spheres=n_spheres)


def camera_rays(s2c, c2w, width, height, seed=0, pixels=None):
    """Primary rays exactly as __raygen__main builds them (main.cu:53-75), vectorised in fp32 numpy;
    jitter from the reference RNG Init(4, pixel, seed)."""
    px = np.arange(width * height, dtype=np.uint32) if pixels is None else np.asarray(pixels, np.uint32)
    v0, v1, s0 = px.copy(), np.full_like(px, seed), np.uint32(0)
    with np.errstate(over="ignore"):
        for _ in range(4):
            s0 = np.uint32(s0 + np.uint32(0x9E3779B9))
            v0 = v0 + ((((v1 << np.uint32(4)) + np.uint32(0xA341316C)) ^ (v1 + s0) ^ ((v1 >> np.uint32(5)) + np.uint32(0xC8013EA4))))
            v1 = v1 + ((((v0 << np.uint32(4)) + np.uint32(0xAD90777D)) ^ (v0 + s0) ^ ((v0 >> np.uint32(5)) + np.uint32(0x7E95761E))))
        st = v0
        st = st * np.uint32(1664525) + np.uint32(1013904223)
        jx = (st & np.uint32(0xFFFFFF)).astype(np.float32) / np.float32(1 << 24)
        st = st * np.uint32(1664525) + np.uint32(1013904223)
        jy = (st & np.uint32(0xFFFFFF)).astype(np.float32) / np.float32(1 << 24)
    x, y = (px % np.uint32(width)).astype(np.float32), (px // np.uint32(width)).astype(np.float32)
    sx, sy = (x + jx) / np.float32(width), (y + jy) / np.float32(height)
    s2c, c2w = np.asarray(s2c, np.float32).reshape(4, 4), np.asarray(c2w, np.float32).reshape(4, 4)
    d = [(s2c[r, 0] * sx + s2c[r, 1] * sy) + s2c[r, 2] * np.float32(0) + s2c[r, 3] * np.float32(1) for r in range(4)]
    inv = np.float32(1) / d[3]
    d = [d[0] * inv, d[1] * inv, d[2] * inv]
    inv = np.float32(1) / np.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
    d = [v * inv for v in d]
    w = [c2w[r, 0] * d[0] + c2w[r, 1] * d[1] + c2w[r, 2] * d[2] for r in range(3)]
    inv = np.float32(1) / np.sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2])
    rays = np.zeros(len(px), capi.RAY_DTYPE)
    rays["ox"], rays["oy"], rays["oz"] = c2w[0, 3], c2w[1, 3], c2w[2, 3]
    rays["dx"], rays["dy"], rays["dz"] = w[0] * inv, w[1] * inv, w[2] * inv
    rays["tmin"], rays["tmax"] = 1e-3, 1e16
    return rays
