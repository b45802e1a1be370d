// pb_geom.cuh — hit -> local shading geometry.  Follows Geometry::GetHitLocalGeometry
// (framework/render/geometry.h:48-96) with the OptiX object/world transforms written out: position
// and texcoord are interpolated in object space and transformed by the instance's 3x4; normals go
// through the inverse transpose; then the two-sided flip (geometry.h:92-96).
#pragma once
#include "pb_scene.cuh"
#include "pb_shade.cuh"

namespace pb {

struct LocalGeo { V3 position, normal; V2 texcoord; };

PB_HD V3 xform_point(const float* m, V3 p) {
    return {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
            m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]};
}
// n'_j = sum_i n_i * inv[i][j]  (optixTransformNormalFromObjectToWorldSpace)
PB_HD V3 xform_normal(const float* inv, V3 n) {
    return {inv[0] * n.x + inv[4] * n.y + inv[8] * n.z, inv[1] * n.x + inv[5] * n.y + inv[9] * n.z,
            inv[2] * n.x + inv[6] * n.y + inv[10] * n.z};
}

PB_HD LocalGeo hit_geometry(const SceneGeom& g, const ObjRec& o, uint32_t prim, float t, float bu, float bv, V3 ray_o, V3 ray_d,
                            bool twosided) {
    LocalGeo r;
    r.texcoord = v2(0.f, 0.f);   // normalised: the reference leaves it stale for meshes without uvs
    if (!(o.flags & OBJ_SPHERE)) {
        const float* T = g.tri + (size_t)kTriShadeFloats * (o.index_offset + (prim - o.prim_offset));
        float f[24];
#if defined(__CUDA_ARCH__)
        // three 256-bit loads (LDG.E.ENL2.256.CONSTANT): one request per 32-byte sector instead of two
#pragma unroll
        for (int k = 0; k < 3; ++k)
            asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=f"(f[8 * k]), "=f"(f[8 * k + 1]), "=f"(f[8 * k + 2]), "=f"(f[8 * k + 3]), "=f"(f[8 * k + 4]),
                           "=f"(f[8 * k + 5]), "=f"(f[8 * k + 6]), "=f"(f[8 * k + 7])
                         : "l"(T + 8 * k));
#else
        for (int k = 0; k < 24; ++k) f[k] = T[k];
#endif
        const V3 p0 = {f[0], f[1], f[2]}, p1 = {f[3], f[4], f[5]}, p2 = {f[6], f[7], f[8]};
        const float w = 1.f - bu - bv;
        r.position = xform_point(o.M, w * p0 + bu * p1 + bv * p2);
        V3 n;
        if (o.flags & OBJ_HAS_NORMALS) n = w * v3(f[9], f[10], f[11]) + bu * v3(f[12], f[13], f[14]) + bv * v3(f[15], f[16], f[17]);
        else n = cross(p1 - p0, p2 - p0);
        r.normal = normalize(xform_normal(o.Minv, n));
        if (o.flags & OBJ_FLIP_NORMALS) r.normal *= -1.f;
        if (o.flags & OBJ_HAS_UV) {
            r.texcoord = w * v2(f[18], f[19]) + bu * v2(f[20], f[21]) + bv * v2(f[22], f[23]);
            if (o.flags & OBJ_FLIP_TEX) r.texcoord.y = 1.f - r.texcoord.y;
        }
    } else {
        r.position = ray_o + t * ray_d;
        const V3 lp = xform_point(o.Minv, r.position);   // sphere centre is the object origin (render_object.cpp:23)
        r.texcoord = sphere_texcoord(normalize(lp - v3(0.f)));
        r.normal = normalize(xform_normal(o.Minv, lp - v3(0.f)));
        if (o.flags & OBJ_FLIP_NORMALS) r.normal *= -1.f;
    }
    if (dot(-ray_d, r.normal) < 0.f && twosided) r.normal = -r.normal;
    return r;
}

}  // namespace pb
