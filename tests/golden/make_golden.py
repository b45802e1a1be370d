#!/usr/bin/env python
"""Generates tests/golden/*.npz from oracle/_ref — the oracle build that compiles the REFERENCE's own
headers (framework/cuda/random.h, render/material/bsdf/*.h, fresnel.h, ggx.h, render/emitter/*.h,
optix/util.h) from /root/reference unmodified — with backend 1 ("reference-headers").  Run in the
dev container only (the reference tree does not exist on the GPU box); the fixtures are committed.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from pupiloptixlab_b200 import capi, scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
RNG_PAIRS = [(7, 0), (0, 0), (1, 0), (65535, 3), (256 * 256 - 1, 15), (123456, 789), (2073599, 255), (0xFFFFFFFF, 0xFFFFFFFF)]


def bsdf_inputs(n, seed):
    rng = np.random.RandomState(seed)
    wo = rng.normal(size=(n, 3))
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    wo[: n // 8, 2] = np.abs(wo[: n // 8, 2]) * 0.02          # grazing
    wo[n // 8: n // 4] = [0, 0, 1]                              # normal incidence
    wo[n // 4: n // 4 + n // 16, 2] *= -1                       # from below
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    seeds = rng.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
    return wo.astype(np.float32), seeds


def golden_materials():
    """(name, material) list: the seven types with the parameters of data/static/material_test.xml plus
    variations that exercise checkerboard textures, nonlinear plastic and the eta < 1 branches."""
    M = capi.make_material
    au = dict(eta=(0.200438, 0.924033, 1.10221), k=(3.91295, 2.45285, 2.14219))
    chk = capi.checker_texture((0.8, 0.2, 0.1), (0.1, 0.3, 0.9), 4, 4)
    return [
        ("diffuse", M(capi.MAT_DIFFUSE, reflectance=(0.7, 0.5, 0.3))),
        ("diffuse_checker", M(capi.MAT_DIFFUSE, reflectance=chk)),
        ("dielectric", M(capi.MAT_DIELECTRIC, int_ior=1.5, ext_ior=1.0)),
        ("dielectric_bk7_tinted", M(capi.MAT_DIELECTRIC, specular_reflectance=(0.9, 0.8, 0.7), specular_transmittance=(0.6, 0.9, 0.8))),
        ("dielectric_inverted", M(capi.MAT_DIELECTRIC, int_ior=1.0, ext_ior=1.33)),
        ("rough_dielectric", M(capi.MAT_ROUGH_DIELECTRIC, int_ior=1.5, ext_ior=1.0, alpha=0.35)),
        ("rough_dielectric_smooth", M(capi.MAT_ROUGH_DIELECTRIC, alpha=0.05, specular_transmittance=(0.9, 0.9, 0.5))),
        ("conductor", M(capi.MAT_CONDUCTOR, specular_reflectance=0.8, **au)),
        ("conductor_mirror", M(capi.MAT_CONDUCTOR)),
        ("rough_conductor", M(capi.MAT_ROUGH_CONDUCTOR, alpha=0.35, specular_reflectance=0.8, **au)),
        ("rough_conductor_checker_alpha", M(capi.MAT_ROUGH_CONDUCTOR, alpha=capi.checker_texture((0.1,) * 3, (0.6,) * 3, 2, 2), **au)),
        ("plastic", M(capi.MAT_PLASTIC, int_ior=1.5, ext_ior=1.0, reflectance=(0.65, 0.3, 0.2))),
        ("plastic_nonlinear", M(capi.MAT_PLASTIC, nonlinear=True, reflectance=chk, specular_reflectance=(0.9, 0.9, 0.8))),
        ("rough_plastic", M(capi.MAT_ROUGH_PLASTIC, int_ior=1.5, ext_ior=1.0, alpha=0.35, reflectance=(0.2, 0.4, 0.65))),
        ("rough_plastic_nonlinear", M(capi.MAT_ROUGH_PLASTIC, nonlinear=True, alpha=0.15, reflectance=(0.5, 0.5, 0.1))),
    ]


def emitter_inputs(n, seed):
    rng = np.random.RandomState(seed)
    hp = rng.uniform(-0.9, 0.9, size=(n, 3)) + [0, 1, 0]
    hn = rng.normal(size=(n, 3))
    hn /= np.linalg.norm(hn, axis=1, keepdims=True)
    hn[: n // 2] = [0, 1, 0]
    xi = rng.rand(n, 2)
    return hp.astype(np.float32), hn.astype(np.float32), xi.astype(np.float32)


RENDERS = [("diffuse", 64, 64, 8, 4), ("materials", 48, 48, 8, 4), ("materials", 40, 24, 2, 2),
           ("textured", 48, 48, 6, 4), ("envmap", 48, 48, 6, 4)]   # bitmap textures / environment map


def generate(ctx_factory, n_bsdf=1500, n_emit=600):
    out = {}
    o = ctx_factory()
    for i, (a, b) in enumerate(RNG_PAIRS):
        seed, fl = o.rng_stream(a, b, 16)
        out[f"rng_seed_{i}"] = np.uint32(seed)
        out[f"rng_floats_{i}"] = fl
    o.scene_begin()
    ids = [o.add_material(m) for _, m in golden_materials()]
    for k, (name, _) in enumerate(golden_materials()):
        wo, seeds = bsdf_inputs(n_bsdf, 100 + k)
        out[f"bsdf_{name}"] = o.bsdf_table(ids[k], wo, seeds)
    o2 = ctx_factory()
    scenes.cornell_box(o2, 64, 64, "materials")      # 2 triangle emitters (Light) + 1 sphere emitter (bulb)
    for ei in range(3):
        hp, hn, xi = emitter_inputs(n_emit, 200 + ei)
        out[f"emitter_{ei}"] = o2.emitter_table(ei, hp, hn, xi)
    for variant, w, h, depth, spp in RENDERS:
        r = ctx_factory()
        scenes.cornell_box(r, w, h, variant)
        s, a, nrm = r.render(w, h, depth, 0, spp)
        key = f"render_{variant}_{w}x{h}_d{depth}_s{spp}"
        out[key + "_sum"], out[key + "_albedo"], out[key + "_normal"] = s, a, nrm
        out[key + "_rays"] = np.array([r.stats()["rays_closest"], r.stats()["rays_any"]], np.uint64)
    return out


def main():
    api = pyoracle.load(prefer_ref=True)
    probe = pyoracle.OracleContext(api)
    if not probe.has_reference():
        raise SystemExit("oracle/_ref is not built (needs /root/reference): run `make -C oracle ref` in the dev container")
    data = generate(lambda: pyoracle.OracleContext(api, backend=1))
    np.savez_compressed(os.path.join(HERE, "reference_headers.npz"), **data)
    print("wrote", os.path.join(HERE, "reference_headers.npz"), len(data), "arrays")


if __name__ == "__main__":
    main()
