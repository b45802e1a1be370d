// stub of framework/util/type.h — only what the binding reads
#pragma once
namespace Pupil::util {
struct Float3 {                     // REF framework/util/type.h:7 | struct Float3 {
    float x, y, z;                  // REF framework/util/type.h:10 | float x, y, z;
};
struct Mat4 {
    float e[16];                    // REF framework/util/type.h:81 | float e[16];
};
}  // namespace Pupil::util
