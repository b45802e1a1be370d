"""The intersection contract against geometry, not against itself.

The reference's ray/triangle and ray/sphere tests live in the closed OptiX 7.5 driver (call sites
example/path_tracer/main.cu:77,158, framework/render/emitter.h:95; sphere module framework/optix/module.cpp:98), so the
oracle DEFINES them and that part of the parity is "unpinned".  What can be checked without OptiX is
that the definition is the geometric closest hit: an independent float64 numpy evaluation (Moeller-Trumbore, unit
sphere under the instance's inverse, nearest root in (tmin, tmax)) must pick the same primitive as the oracle's fp32
brute force, except where float64 itself sees a near tie or an edge graze.
"""
import numpy as np
import pytest

from oracle import pyoracle
from pupiloptixlab_b200 import capi
from tests import util_scenes as U


def f64_closest(pos, idx, sph, rays):
    o = np.stack([rays["ox"], rays["oy"], rays["oz"]], -1).astype(np.float64)
    d = np.stack([rays["dx"], rays["dy"], rays["dz"]], -1).astype(np.float64)
    tmin, tmax = rays["tmin"].astype(np.float64), rays["tmax"].astype(np.float64)
    n = len(rays)
    T = np.full((n, len(idx) + len(sph)), np.inf)
    edge = np.full((n, len(idx) + len(sph)), np.inf)      # distance of (u, v) from the triangle's edges
    if len(idx):
        P = pos.astype(np.float64)
        v0, v1, v2 = P[idx[:, 0]], P[idx[:, 1]], P[idx[:, 2]]
        e1, e2 = v1 - v0, v2 - v0
        p = np.cross(d[:, None, :], e2[None])
        det = np.einsum("tk,ntk->nt", e1, p)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / det
            s = o[:, None, :] - v0[None]
            u = np.einsum("ntk,ntk->nt", s, p) * inv
            q = np.cross(s, e1[None])
            v = np.einsum("nk,ntk->nt", d, q) * inv
            t = np.einsum("tk,ntk->nt", e2, q) * inv
        ok = (det != 0) & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > tmin[:, None]) & (t < tmax[:, None])
        T[:, :len(idx)] = np.where(ok, t, np.inf)
        edge[:, :len(idx)] = np.minimum(np.minimum(np.abs(u), np.abs(v)), np.abs(1 - u - v))
    for k, M in enumerate(sph):
        A = np.vstack([M.astype(np.float64), [0, 0, 0, 1]])
        Ai = np.linalg.inv(A)
        oo = o @ Ai[:3, :3].T + Ai[:3, 3]
        dd = d @ Ai[:3, :3].T
        a = (dd * dd).sum(-1); b = (oo * dd).sum(-1); c = (oo * oo).sum(-1) - 1.0
        disc = b * b - a * c
        with np.errstate(invalid="ignore", divide="ignore"):
            sq = np.sqrt(np.maximum(disc, 0))
            t0, t1 = (-b - sq) / a, (-b + sq) / a
        in0 = (disc >= 0) & (t0 > tmin) & (t0 < tmax)
        in1 = (disc >= 0) & (t1 > tmin) & (t1 < tmax)
        T[:, len(idx) + k] = np.where(in0, t0, np.where(in1, t1, np.inf))
        with np.errstate(invalid="ignore"):
            edge[:, len(idx) + k] = np.abs(disc) / np.maximum(b * b, 1e-300)   # grazing: discriminant near zero
    return T, edge


@pytest.mark.parametrize("n_tri,n_sph,seed", [(300, 0, 11), (0, 40, 12), (400, 30, 13)])
def test_oracle_brute_force_is_the_geometric_closest_hit(oracle_api, n_tri, n_sph, seed):
    pos, idx, sph = U.random_soup(n_tri, n_sph, seed)
    o = pyoracle.OracleContext(oracle_api)
    U.oracle_scene(o, pos, idx, sph)
    rays = U.random_rays(6000, seed + 1)
    rays["tmax"][::4] = 2.0
    got = o.trace_closest_brute(rays)
    T, edge = f64_closest(pos, idx, sph, rays)
    want = np.where(np.isfinite(T.min(1)), T.argmin(1), -1)
    got_prim = np.where(got["prim"] == capi.MISS, -1, got["prim"].astype(np.int64))
    same = got_prim == want
    assert same.mean() > 0.995, same.mean()
    assert (want >= 0).sum() > 500
    # where both agree on the primitive, t agrees to fp32 accuracy
    both = same & (want >= 0)
    t64 = T[np.arange(len(rays)), np.maximum(want, 0)]
    rel = np.abs(got["t"][both].astype(np.float64) - t64[both]) / np.maximum(np.abs(t64[both]), 1e-3)
    assert rel.max() < 2e-4, rel.max()
    # a disagreement must be explainable in float64: a near tie in t, an edge / silhouette graze, or a hit at the very ends
    # of (tmin, tmax)
    for r in np.nonzero(~same)[0]:
        cands = [p for p in (got_prim[r], want[r]) if p >= 0]
        graze = any(edge[r, p] < 1e-4 for p in cands)
        tie = len(cands) == 2 and np.isfinite(T[r, cands]).all() and abs(T[r, cands[0]] - T[r, cands[1]]) < 1e-4 * max(1.0, abs(T[r, cands[1]]))
        tv = [T[r, p] for p in cands if np.isfinite(T[r, p])] + ([float(got["t"][r])] if got_prim[r] >= 0 else [])
        at_end = any(abs(t - rays["tmin"][r]) < 1e-5 or abs(t - rays["tmax"][r]) < 1e-4 * rays["tmax"][r] for t in tv)
        assert graze or tie or at_end, (r, got_prim[r], want[r], [T[r, p] for p in cands], [edge[r, p] for p in cands])
