#pragma once
#include <cstddef>
#include "transform.h"
#include "type.h"
namespace Pupil::util {
enum class ETextureAddressMode : unsigned int { Wrap = 0, Clamp = 1, Mirror = 2, Border = 3 };   // REF framework/util/texture.h:11 | Wrap = 0,
enum class ETextureFilterMode : unsigned int { Point = 0, Linear = 1 };                           // REF framework/util/texture.h:18 | Point = 0,
enum class ETextureType : unsigned int {
    RGB = 0,                        // REF framework/util/texture.h:23 | RGB = 0,
    Bitmap,                         // REF framework/util/texture.h:24 | Bitmap,
    Checkerboard                    // REF framework/util/texture.h:25 | Checkerboard
};
struct RGBTexture {
    Float3 color{0.f, 0.f, 0.f};    // REF framework/util/texture.h:29 | Float3 color{ 0.f, 0.f, 0.f };
};
struct CheckerboardTexture {
    Float3 patch1{0.4f, 0.4f, 0.4f};   // REF framework/util/texture.h:33 | Float3 patch1{ 0.4f, 0.4f, 0.4f };
    Float3 patch2{0.2f, 0.2f, 0.2f};   // REF framework/util/texture.h:34 | Float3 patch2{ 0.2f, 0.2f, 0.2f };
};
struct BitmapTexture {
    size_t w = 0;                   // REF framework/util/texture.h:38 | size_t w = 0;
    size_t h = 0;                   // REF framework/util/texture.h:39 | size_t h = 0;
    float* data = nullptr;          // REF framework/util/texture.h:40 | float *data = nullptr;
    ETextureAddressMode address_mode = ETextureAddressMode::Wrap;   // REF framework/util/texture.h:42 | ETextureAddressMode address_mode = ETextureAddressMode::Wrap;
    ETextureFilterMode filter_mode = ETextureFilterMode::Linear;    // REF framework/util/texture.h:43 | ETextureFilterMode filter_mode = ETextureFilterMode::Linear;
};
struct Texture {
    ETextureType type = ETextureType::RGB;   // REF framework/util/texture.h:58 | ETextureType type = ETextureType::RGB;
    union {
        RGBTexture rgb;                      // REF framework/util/texture.h:60 | RGBTexture rgb;
        BitmapTexture bitmap;                // REF framework/util/texture.h:61 | BitmapTexture bitmap;
        CheckerboardTexture checkerboard;    // REF framework/util/texture.h:62 | CheckerboardTexture checkerboard;
    };
    Transform transform{};                   // REF framework/util/texture.h:64 | Transform transform{};
    Texture() noexcept : rgb(RGBTexture{}) {}
};
}  // namespace Pupil::util
