#pragma once
#include "type.h"
namespace Pupil::util {
struct Transform {
    Mat4 matrix;                    // REF framework/util/transform.h:7 | Mat4 matrix;
};
}  // namespace Pupil::util
