#pragma once
#include "util/texture.h"
namespace Pupil {
enum class EMatType : unsigned int {         // REF framework/render/material/predefine.h:15 | enum class EMatType : unsigned int {
    Unknown = 0,                             // REF framework/render/material/predefine.h:16 | Unknown = 0,
    Diffuse,                                 // REF framework/decl/material_decl.inl:4 | PUPIL_MATERIAL_TYPE_DEFINE(Diffuse)
    Dielectric,                              // REF framework/decl/material_decl.inl:5 | PUPIL_MATERIAL_TYPE_DEFINE(Dielectric)
    RoughDielectric,                         // REF framework/decl/material_decl.inl:6 | PUPIL_MATERIAL_TYPE_DEFINE(RoughDielectric)
    Conductor,                               // REF framework/decl/material_decl.inl:7 | PUPIL_MATERIAL_TYPE_DEFINE(Conductor)
    RoughConductor,                          // REF framework/decl/material_decl.inl:8 | PUPIL_MATERIAL_TYPE_DEFINE(RoughConductor)
    Plastic,                                 // REF framework/decl/material_decl.inl:9 | PUPIL_MATERIAL_TYPE_DEFINE(Plastic)
    RoughPlastic,                            // REF framework/decl/material_decl.inl:10 | PUPIL_MATERIAL_TYPE_DEFINE(RoughPlastic)
    Twosided,
    Count
};
}  // namespace Pupil
namespace Pupil::resource {
struct Diffuse {
    util::Texture reflectance;               // REF framework/resource/material.h:17 | util::Texture reflectance;
};
struct Dielectric {
    float int_ior;                           // REF framework/resource/material.h:21 | float int_ior;
    float ext_ior;                           // REF framework/resource/material.h:22 | float ext_ior;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:23 | util::Texture specular_reflectance;
    util::Texture specular_transmittance;    // REF framework/resource/material.h:24 | util::Texture specular_transmittance;
};
struct RoughDielectric {
    float int_ior;                           // REF framework/resource/material.h:28 | float int_ior;
    float ext_ior;                           // REF framework/resource/material.h:29 | float ext_ior;
    util::Texture alpha;                     // REF framework/resource/material.h:30 | util::Texture alpha;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:31 | util::Texture specular_reflectance;
    util::Texture specular_transmittance;    // REF framework/resource/material.h:32 | util::Texture specular_transmittance;
};
struct Conductor {
    util::Texture eta;                       // REF framework/resource/material.h:36 | util::Texture eta;
    util::Texture k;                         // REF framework/resource/material.h:37 | util::Texture k;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:38 | util::Texture specular_reflectance;
};
struct RoughConductor {
    util::Texture alpha;                     // REF framework/resource/material.h:42 | util::Texture alpha;
    util::Texture eta;                       // REF framework/resource/material.h:43 | util::Texture eta;
    util::Texture k;                         // REF framework/resource/material.h:44 | util::Texture k;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:45 | util::Texture specular_reflectance;
};
struct Plastic {
    float int_ior;                           // REF framework/resource/material.h:49 | float int_ior;
    float ext_ior;                           // REF framework/resource/material.h:50 | float ext_ior;
    bool nonlinear;                          // REF framework/resource/material.h:51 | bool nonlinear;
    util::Texture diffuse_reflectance;       // REF framework/resource/material.h:52 | util::Texture diffuse_reflectance;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:53 | util::Texture specular_reflectance;
};
struct RoughPlastic {
    float int_ior;                           // REF framework/resource/material.h:57 | float int_ior;
    float ext_ior;                           // REF framework/resource/material.h:58 | float ext_ior;
    bool nonlinear;                          // REF framework/resource/material.h:59 | bool nonlinear;
    util::Texture alpha;                     // REF framework/resource/material.h:60 | util::Texture alpha;
    util::Texture diffuse_reflectance;       // REF framework/resource/material.h:61 | util::Texture diffuse_reflectance;
    util::Texture specular_reflectance;      // REF framework/resource/material.h:62 | util::Texture specular_reflectance;
};
struct Material {
    EMatType type = EMatType::Unknown;       // REF framework/resource/material.h:69 | EMatType type = EMatType::Unknown;
    bool twosided = false;                   // REF framework/resource/material.h:70 | bool twosided = false;
    union {
        Diffuse diffuse{};                   // REF framework/resource/material.h:73 | Diffuse diffuse{};
        Dielectric dielectric;               // REF framework/resource/material.h:74 | Dielectric dielectric;
        RoughDielectric rough_dielectric;    // REF framework/resource/material.h:75 | RoughDielectric rough_dielectric;
        Conductor conductor;                 // REF framework/resource/material.h:76 | Conductor conductor;
        RoughConductor rough_conductor;      // REF framework/resource/material.h:77 | RoughConductor rough_conductor;
        Plastic plastic;                     // REF framework/resource/material.h:78 | Plastic plastic;
        RoughPlastic rough_plastic;          // REF framework/resource/material.h:79 | RoughPlastic rough_plastic;
    };
    Material() noexcept {}
};
}  // namespace Pupil::resource
