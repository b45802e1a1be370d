// pb_types.h — device-side data layout of the B200 path tracer and the fp32 helper algebra.
//
// HBM layout (all arrays are plain cudaMalloc'd, 16-byte aligned, read through 128-bit loads):
//   Node8   80 B   compressed 8-wide BVH node (5 x uint4), see pb_traverse.cuh
//   PrimRec 48 B   world-space primitive in leaf order (3 x float4): triangle or sphere
//   ObjRec  128 B  per render object: transforms, material, emitter offset, mesh buffer offsets
//   MatRec  320 B  material with up to 6 inline textures (RGB / checkerboard inline, bitmap handle)
//   EmitRec 192 B  one per emissive triangle / sphere (+1 env)
// Path state is SoA float4 arrays indexed by path slot (pb_wavefront.cu).
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define PB_HD __host__ __device__ __forceinline__
#define PB_D __device__ __forceinline__
#else
#define PB_HD inline
#define PB_D inline
#endif

namespace pb {

// ---- exactly-rounded scalar ops: never contracted, identical on host and device -----------------
#if defined(__CUDA_ARCH__)
PB_HD float f_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
PB_HD float f_mul(float a, float b) { return __fmul_rn(a, b); }
PB_HD float f_add(float a, float b) { return __fadd_rn(a, b); }
PB_HD float f_sub(float a, float b) { return __fsub_rn(a, b); }
PB_HD float f_div(float a, float b) { return __fdiv_rn(a, b); }
PB_HD float f_sqrt(float a) { return __fsqrt_rn(a); }
#else
PB_HD float f_fma(float a, float b, float c) { return fmaf(a, b, c); }
PB_HD float f_mul(float a, float b) { return a * b; }
PB_HD float f_add(float a, float b) { return a + b; }
PB_HD float f_sub(float a, float b) { return a - b; }
PB_HD float f_div(float a, float b) { return a / b; }
PB_HD float f_sqrt(float a) { return sqrtf(a); }
#endif

PB_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
PB_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// ---- small vector algebra.  Operation order follows the reference's vec_math.h (dot :462,
// normalize :477, v/s == v*(1/s) :425) because image parity depends on it; the translation unit is
// compiled with -fmad=false so a*b+c is never fused.
struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

PB_HD V3 v3(float s) { return {s, s, s}; }
PB_HD V3 v3(float x, float y, float z) { return {x, y, z}; }
PB_HD V2 v2(float x, float y) { return {x, y}; }
PB_HD V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
PB_HD V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
PB_HD V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
PB_HD V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
PB_HD V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
PB_HD V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
PB_HD V3 operator/(V3 a, float s) { float inv = 1.0f / s; return a * inv; }
PB_HD V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
PB_HD V3 operator-(float s, V3 a) { return {s - a.x, s - a.y, s - a.z}; }
PB_HD V3& operator+=(V3& a, V3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
PB_HD V3& operator*=(V3& a, V3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
PB_HD V3& operator*=(V3& a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
PB_HD V3& operator/=(V3& a, float s) { float inv = 1.0f / s; a *= inv; return a; }
PB_HD V2 operator+(V2 a, V2 b) { return {a.x + b.x, a.y + b.y}; }
PB_HD V2 operator*(V2 a, float s) { return {a.x * s, a.y * s}; }
PB_HD V2 operator*(float s, V2 a) { return {a.x * s, a.y * s}; }
PB_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PB_HD float dot(V4 a, V4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
PB_HD V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
PB_HD float length(V3 v) { return sqrtf(dot(v, v)); }
PB_HD V3 normalize(V3 v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }
PB_HD V4 normalize(V4 v) { float inv = 1.0f / sqrtf(dot(v, v)); return {v.x * inv, v.y * inv, v.z * inv, v.w * inv}; }

// ---- compressed 8-wide BVH node (Ylitie/Karras/Laine 2017 layout), 80 bytes -----------------------
struct alignas(16) Node8 {
    float px, py, pz;          // quantisation origin
    uint8_t ex, ey, ez, imask; // per-axis exponent (biased, scale = 2^(e-127) as float bits e<<23), inner mask
    uint32_t child_base;       // index of the first inner child (children are contiguous, rank = popc)
    uint32_t prim_base;        // index of the first primitive referenced by leaf children
    uint32_t pmask;            // leaf slots: 3 bits per slot (bits 3s..3s+2), unary primitive count (1, 3, 7); inner / empty: 0.
                               // Primitive j of the node's range sits behind the j-th set bit (slot order).
    uint32_t spare;
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

// ---- primitive in leaf order, 48 bytes ----------------------------------------------------------
// triangle: a = (v0, prim id), b = (v1, object id), c = (v2, kind = 0)
// sphere  : a = (centre, prim id), b = (half extents of the world box, object id), c = (0,0,0, kind = 1)
// sphere whose object->world linear part is s * identity (what ShapeLoader<_sphere> produces, shape.cpp:106-125, unless
//           the XML adds a rotation / non-uniform scale): a = (centre, prim id), b = (|s| = every half extent, 1/s = the
//           diagonal of world->object, 0, object id), c = (translation column of world->object, kind = 2).  The record
//           then holds everything the unit-sphere test needs (pb_intersect.cuh sphere_test_uniform): no dependent fetch
//           of the object's inverse matrix, and the arithmetic is the general path's with its zero terms dropped,
//           i.e. bit-identical results (the oracle keeps the general form only).
struct alignas(16) PrimRec { V4 a, b, c; };
static_assert(sizeof(PrimRec) == 48, "PrimRec must be 48 bytes");
enum { PRIM_TRI = 0u, PRIM_SPHERE = 1u, PRIM_SPHERE_UNIFORM = 2u };

// ---- textures / materials / objects / emitters --------------------------------------------------
struct TexRec {          // 64 B; mirror of cuda::Texture (framework/cuda/texture.h:9-31)
    uint32_t type;       // 0 rgb, 1 bitmap, 2 checkerboard
    float a[3];          // rgb | patch1 | bitmap: bits of (address mode, filter mode, height)
    float b[3];          // patch2 | bitmap: b[0..1] = the 64 bits of the device pointer to the RGBA float texels
    uint32_t bitmap_w;   // bitmap: width
    float r0[4];
    float r1[4];
};
// bitmap view of a TexRec (the reference keeps a cudaTextureObject_t in the same union, texture.h:12-19)
struct TexBitmapView { const float* texels; uint32_t w, h, address_mode, filter_mode; };
PB_HD TexBitmapView tex_bitmap(const TexRec& t) {
    TexBitmapView v;
    v.texels = reinterpret_cast<const float*>((unsigned long long)f2u(t.b[0]) | ((unsigned long long)f2u(t.b[1]) << 32));
    v.w = t.bitmap_w; v.h = f2u(t.a[2]); v.address_mode = f2u(t.a[0]); v.filter_mode = f2u(t.a[1]);
    return v;
}

enum { TEX_REFLECTANCE = 0, TEX_SPEC_REFL, TEX_SPEC_TRANS, TEX_ALPHA, TEX_ETA, TEX_K, TEX_COUNT };

struct alignas(16) MatRec {
    uint32_t type;       // EMatType 1..7
    uint32_t twosided;
    float eta;           // int_ior / ext_ior
    float int_ior_over_ext_dielectric;  // dielectric.h:50 recomputes int/ext: same value, kept explicit
    float int_fdr;
    float specular_sampling_weight;
    uint32_t nonlinear;
    uint32_t _pad;
    TexRec tex[TEX_COUNT];
};

enum { OBJ_SPHERE = 1u, OBJ_FLIP_NORMALS = 2u, OBJ_FLIP_TEX = 4u, OBJ_HAS_NORMALS = 8u, OBJ_HAS_UV = 16u };
struct alignas(16) ObjRec {   // 128 B
    float M[12];              // object -> world rows
    float Minv[12];           // world -> object rows
    uint32_t flags;
    uint32_t material;
    int32_t emitter_offset;   // -1 if not an emitter
    uint32_t prim_offset;     // global id of its first primitive
    uint32_t vert_offset;     // into the global vertex attribute arrays
    uint32_t index_offset;    // into the global index array (in triangles)
    uint32_t n_prims;
    uint32_t _pad;
};
static_assert(sizeof(ObjRec) == 128, "ObjRec must be 128 bytes");

enum { EM_NONE = 0u, EM_TRI = 1u, EM_SPHERE = 2u, EM_CONST_ENV = 3u, EM_ENV_MAP = 4u };
struct alignas(16) EmitRec {
    uint32_t type;
    float weight, select_probability, area;
    TexRec radiance;
    float p0[3], p1[3], p2[3];   // tri: world positions | sphere: p0 = centre, p1[0] = radius | env: p0 = colour, p1 = centre
    float n0[3], n1[3], n2[3];
    float t0[2], t1[2], t2[2];
    // env map
    uint32_t map_w, map_h, n_row_cdf, _pad;
    const float* row_cdf;
    const float* col_cdf;
    const float* row_weight;
    float to_world[9], to_local[9];
    float normalization, scale;
};

struct Camera { float s2c[16]; float c2w[16]; };

}  // namespace pb
