// pb_build.cuh — per-thread bodies of the device BVH build (kernels in pb_build.cu wrap them).
//
// Replaces optixAccelBuild for GAS + IAS (framework/world/gas_manager.cpp:61-185,
// framework/world/ias_manager.cpp:29-151): instances are flattened to world space and ONE
// hierarchy is built over all primitives:
//   1. padded boxes B_out + 63-bit Morton keys of box centres
//   2. radix sort (CUB, in pb_build.cu)
//   3. binary radix tree (Karras 2012) + bottom-up boxes/counts with one atomic flag per node
//   4. level-by-level collapse into 8-wide nodes: greedy opening by surface area, octant slot
//      assignment, conservative 8-bit quantisation, primitives copied into leaf order
// Tree shape never influences results (pb_intersect.cuh), only speed.
//
// Every function is PB_HD so tests/hostsim can run the same code on the CPU.
#pragma once
#include "pb_traverse.cuh"

namespace pb {

constexpr uint32_t kLeafFlag = 0x80000000u;
constexpr uint32_t kNoRef = 0xFFFFFFFFu;
constexpr uint32_t kLeafMax = 3u;   // primitives per leaf slot (unary count in 3 bits)

struct BuildArrays {
    // inputs
    const PrimRec* prims_in;    // insertion order (global prim id = index)
    uint32_t n;
    float delta;
    V3 scene_lo, scene_hi;      // exact scene box
    // per primitive (insertion order)
    V3* blo; V3* bhi;           // B_out boxes
    unsigned long long* keys;   // Morton keys (sorted in place together with `order`)
    uint32_t* order;            // sorted position -> prim id
    // binary tree (n-1 internal nodes)
    uint32_t* left; uint32_t* right;      // child refs (leaf: sorted position | kLeafFlag)
    uint32_t* parent_int; uint32_t* parent_leaf;
    uint32_t* first; uint32_t* last;      // sorted range of each internal node
    V3* ilo; V3* ihi;                     // internal boxes
    uint32_t* flag;                       // arrival counters
    uint32_t leaf_max;                    // primitives per leaf slot, 1..3 (kLeafMax)
    uint32_t bottom_max;                  // subtrees with at most this many primitives become ONE wide node (<= 24)
};

PB_HD unsigned long long expand21(uint32_t v) {  // spread 21 bits to every third bit
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// step 1: padded box + Morton key of primitive i
PB_HD void build_prim(const BuildArrays& b, uint32_t i) {
    V3 lo, hi;
    prim_bbox(b.prims_in[i], lo, hi);
    pad_box(lo, hi, 2.0f * b.delta);
    b.blo[i] = lo; b.bhi[i] = hi;
    V3 ext = b.scene_hi - b.scene_lo;
    V3 c = (lo + hi) * 0.5f - b.scene_lo;
    float fx = ext.x > 0.f ? c.x / ext.x : 0.f, fy = ext.y > 0.f ? c.y / ext.y : 0.f, fz = ext.z > 0.f ? c.z / ext.z : 0.f;
    fx = fminf(fmaxf(fx, 0.f), 1.f); fy = fminf(fmaxf(fy, 0.f), 1.f); fz = fminf(fmaxf(fz, 0.f), 1.f);
    uint32_t qx = (uint32_t)fminf(fx * 2097152.f, 2097151.f), qy = (uint32_t)fminf(fy * 2097152.f, 2097151.f),
             qz = (uint32_t)fminf(fz * 2097152.f, 2097151.f);
    b.keys[i] = (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz);
    b.order[i] = i;
}

PB_HD int clz64(unsigned long long x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
PB_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
// common-prefix length of sorted keys i and j (index tiebreak), -1 out of range
PB_HD int lcp(const BuildArrays& b, int i, int j) {
    if (j < 0 || j >= (int)b.n) return -1;
    unsigned long long a = b.keys[i], c = b.keys[j];
    if (a == c) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz64(a ^ c);
}

// step 3a: internal node i of the binary radix tree (Karras 2012, fig. 4)
PB_HD void build_radix_node(const BuildArrays& b, uint32_t iu) {
    const int i = (int)iu;
    const int d = (lcp(b, i, i + 1) - lcp(b, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lcp(b, i, i - d);
    int lmax = 2;
    while (lcp(b, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lcp(b, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lcp(b, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (lcp(b, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    if (lo == gamma) { b.left[i] = (uint32_t)gamma | kLeafFlag; b.parent_leaf[gamma] = iu; }
    else { b.left[i] = (uint32_t)gamma; b.parent_int[gamma] = iu; }
    if (hi == gamma + 1) { b.right[i] = (uint32_t)(gamma + 1) | kLeafFlag; b.parent_leaf[gamma + 1] = iu; }
    else { b.right[i] = (uint32_t)(gamma + 1); b.parent_int[gamma + 1] = iu; }
    b.first[i] = (uint32_t)lo; b.last[i] = (uint32_t)hi;
    if (i == 0) b.parent_int[0] = kNoRef;
}

PB_HD void ref_box(const BuildArrays& b, uint32_t ref, V3& lo, V3& hi) {
    if (ref & kLeafFlag) { uint32_t p = b.order[ref & ~kLeafFlag]; lo = b.blo[p]; hi = b.bhi[p]; }
    else { lo = b.ilo[ref]; hi = b.ihi[ref]; }
}
PB_HD uint32_t ref_count(const BuildArrays& b, uint32_t ref) {
    return (ref & kLeafFlag) ? 1u : b.last[ref] - b.first[ref] + 1u;
}

// step 3b: bottom-up boxes.  Called once per leaf (sorted position); the second arrival at a node
// computes its box and climbs.  `arrive` must return the previous value of an atomic increment.
template<class Arrive>
PB_HD void build_refit_from_leaf(const BuildArrays& b, uint32_t leaf, Arrive arrive) {
    uint32_t node = b.parent_leaf[leaf];
    while (node != kNoRef) {
        if (arrive(&b.flag[node]) == 0u) return;  // first arrival: the sibling subtree is not done yet
        V3 alo, ahi, clo, chi;
        ref_box(b, b.left[node], alo, ahi);
        ref_box(b, b.right[node], clo, chi);
        b.ilo[node] = {fminf(alo.x, clo.x), fminf(alo.y, clo.y), fminf(alo.z, clo.z)};
        b.ihi[node] = {fmaxf(ahi.x, chi.x), fmaxf(ahi.y, chi.y), fmaxf(ahi.z, chi.z)};
        node = b.parent_int[node];
    }
}

}  // namespace pb
