#pragma once
#include "util/texture.h"
#include "util/type.h"
namespace Pupil::resource {
enum class EEmitterType {
    Unknown,
    Area,                                    // REF framework/resource/emitter.h:9 | Area,
    Point,
    ConstEnv,                                // REF framework/resource/emitter.h:11 | ConstEnv,
    EnvMap,                                  // REF framework/resource/emitter.h:12 | EnvMap,
    DistDir
};
struct AreaEmitter {
    util::Texture radiance;                  // REF framework/resource/emitter.h:17 | util::Texture radiance;
};
struct ConstEnv {
    util::Float3 radiance;                   // REF framework/resource/emitter.h:26 | util::Float3 radiance;
};
struct EnvMap {
    util::Texture radiance;                  // REF framework/resource/emitter.h:30 | util::Texture radiance;
    float scale;                             // REF framework/resource/emitter.h:31 | float scale;
    util::Transform transform{};             // REF framework/resource/emitter.h:33 | util::Transform transform{};
};
struct Emitter {
    EEmitterType type = EEmitterType::Unknown;   // REF framework/resource/emitter.h:42 | EEmitterType type = EEmitterType::Unknown;
    union {
        AreaEmitter area;                    // REF framework/resource/emitter.h:44 | AreaEmitter area;
        ConstEnv const_env;                  // REF framework/resource/emitter.h:46 | ConstEnv const_env;
        EnvMap env_map;                      // REF framework/resource/emitter.h:47 | EnvMap env_map;
    };
    Emitter() noexcept {}
};
}  // namespace Pupil::resource
