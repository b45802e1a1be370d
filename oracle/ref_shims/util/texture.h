// ORACLE shim (test infrastructure): the part of the reference's framework/util/texture.h that
// framework/cuda/texture.h needs on the host — the ETextureType enum (util/texture.h:22-26) and an
// opaque util::Texture for the CudaTextureManager declaration.  The real header pulls DirectXMath.
#pragma once
#include <cstddef>
namespace Pupil::util {
enum class ETextureType : unsigned int { RGB = 0, Bitmap, Checkerboard };
struct Texture { ETextureType type = ETextureType::RGB; };
}  // namespace Pupil::util
