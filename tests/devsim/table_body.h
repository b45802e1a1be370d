// tests/devsim/table_body.h — TEST HARNESS ONLY.  One entry of the per-BSDF known-answer table, computed by the
// PRODUCT's shading headers (pupiloptixlab_b200/csrc/pb_bsdf.cuh, pb_shade.cuh).  Same inputs and the same 16-float
// record as the oracle's orc_bsdf_table (oracle/oracle.cpp), whose reference-header output is committed in
// tests/golden/reference_headers.npz:  material at uv = (0.25, 0.75); Sample(wo, rng seed) -> wi, f, pdf, lobe, rng
// state after; then Eval(wo, wi) -> f, pdf.
// Compiled twice: by g++ into tests/hostsim/libhostsim.so (CPU, bit-exact against the golden table apart from nothing:
// same glibc) and by nvcc into tests/devsim/libdevsim.so (the same code on the GPU, libdevice sin/cos).
#pragma once
#include "../../pupiloptixlab_b200/csrc/pb_bsdf.cuh"

namespace pb {

template<int MAT>
PB_HD void bsdf_table_entry(const MatRec& m, const float* wo, uint32_t seed, float* o) {
    const BsdfLocal b = bsdf_local<MAT>(m, v2(0.25f, 0.75f));
    Rng rng; rng.s = seed;
    BsdfRec r;
    r.wo = {wo[0], wo[1], wo[2]};
    bsdf_sample<MAT>(b, r, rng);
    BsdfRec e;
    e.wo = r.wo; e.wi = r.wi;
    bsdf_eval<MAT>(b, e);
    o[0] = r.wi.x; o[1] = r.wi.y; o[2] = r.wi.z; o[3] = r.f.x; o[4] = r.f.y; o[5] = r.f.z; o[6] = r.pdf;
    o[7] = u2f(r.lobe); o[8] = u2f(rng.s);
    o[9] = e.f.x; o[10] = e.f.y; o[11] = e.f.z; o[12] = e.pdf; o[13] = o[14] = o[15] = 0.f;
}

PB_HD void bsdf_table_entry_any(const MatRec& m, const float* wo, uint32_t seed, float* o) {
    switch (m.type) {
        case MAT_DIFFUSE: bsdf_table_entry<MAT_DIFFUSE>(m, wo, seed, o); break;
        case MAT_DIELECTRIC: bsdf_table_entry<MAT_DIELECTRIC>(m, wo, seed, o); break;
        case MAT_ROUGH_DIELECTRIC: bsdf_table_entry<MAT_ROUGH_DIELECTRIC>(m, wo, seed, o); break;
        case MAT_CONDUCTOR: bsdf_table_entry<MAT_CONDUCTOR>(m, wo, seed, o); break;
        case MAT_ROUGH_CONDUCTOR: bsdf_table_entry<MAT_ROUGH_CONDUCTOR>(m, wo, seed, o); break;
        case MAT_PLASTIC: bsdf_table_entry<MAT_PLASTIC>(m, wo, seed, o); break;
        case MAT_ROUGH_PLASTIC: bsdf_table_entry<MAT_ROUGH_PLASTIC>(m, wo, seed, o); break;
        default: for (int k = 0; k < 16; ++k) o[k] = 0.f;
    }
}

}  // namespace pb
