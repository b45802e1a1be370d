"""Shared helpers for tests: random geometry and rays."""
import numpy as np

from pupiloptixlab_b200 import capi


def random_soup(n_tri, n_sph, seed, axis_aligned=False, extent=2.0):
    rng = np.random.RandomState(seed)
    c = rng.uniform(-extent, extent, size=(n_tri, 1, 3))
    tri = (c + rng.normal(scale=0.15 * extent, size=(n_tri, 3, 3))).astype(np.float32)
    if axis_aligned:  # flatten every triangle into an axis plane at a grid coordinate (ties, flat boxes)
        ax = rng.randint(0, 3, size=n_tri)
        plane = rng.randint(-4, 5, size=n_tri).astype(np.float32) * 0.25
        for k in range(3):
            tri[ax == k, :, k] = plane[ax == k, None]
    pos = tri.reshape(-1, 3)
    idx = np.arange(3 * n_tri, dtype=np.uint32).reshape(-1, 3)
    sph = np.zeros((n_sph, 3, 4), np.float32)
    for s in range(n_sph):
        A = rng.normal(size=(3, 3)) * 0.2 + np.eye(3) * rng.uniform(0.05, 0.4)
        if s % 2 == 0:
            A = np.eye(3) * rng.uniform(0.05, 0.4)
        sph[s, :, :3] = A
        sph[s, :, 3] = rng.uniform(-extent, extent, size=3)
    return pos, idx, sph


def random_rays(n, seed, extent=2.5, tmax=1e16, from_points=None):
    rng = np.random.RandomState(seed)
    rays = np.zeros(n, capi.RAY_DTYPE)
    o = rng.uniform(-extent, extent, size=(n, 3)) if from_points is None else from_points[rng.randint(0, len(from_points), n)]
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = n // 10  # a share of exactly axis-parallel directions (zero components)
    d[:k] = 0
    d[np.arange(k), rng.randint(0, 3, k)] = rng.choice([-1.0, 1.0], k)
    rays["ox"], rays["oy"], rays["oz"] = o.T.astype(np.float32)
    rays["dx"], rays["dy"], rays["dz"] = d.T.astype(np.float32)
    rays["tmin"] = 1e-3
    rays["tmax"] = tmax
    return rays


def oracle_scene(ctx, pos, idx, sph):
    ctx.scene_begin()
    m = ctx.add_material(capi.make_material(capi.MAT_DIFFUSE))
    ident = np.eye(4, dtype=np.float32)
    if len(idx):
        ctx.add_trimesh(pos, None, None, idx, ident, 0, m)
    for s in sph:
        ctx.add_sphere(np.vstack([s, [0, 0, 0, 1]]).astype(np.float32), 0, m)
    ctx.set_camera(np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32))
    ctx.commit()
