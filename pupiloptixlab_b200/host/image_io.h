// image_io.h — image files for bitmap textures / environment maps and for saving renders.
//
// Replaces BitmapTexture::Load / Save of the reference (framework/util/texture.cpp:13-180), which wraps
// stb_image, stb_image_write and tinyexr (third-party, vendored there).  Written from the file-format
// specifications; zlib (system library) does the inflate / deflate.
//
//   load_image : by extension / magic.  8-bit files (PNG, BMP, TGA, PPM/PGM) go through the reference's
//                decode pow(x / 255, 2.2) (texture.cpp:110-115, alpha x / 255 or 1); float files (Radiance
//                .hdr, OpenEXR, PFM) are taken as they are (texture.cpp:95-106,130-147).  Result = RGBA float,
//                row 0 at the top.  JPEG is recognised and refused with a clear message (no decoder here).
//   save_*     : .hdr and .exr follow the reference's writers (rows flipped vertically, texture.cpp:14,36-44;
//                EXR channels B, G, R as 32-bit float); .pfm and 8-bit .png are additions for headless use.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace pupil::host {

struct Image {
    int w = 0, h = 0;
    std::vector<float> rgba;   // w * h * 4
};

bool load_image(const std::string& path, Image& out, std::string& err);
bool load_image_memory(const unsigned char* data, size_t n, const std::string& name_hint, Image& out, std::string& err);

bool save_hdr(const std::string& path, const float* rgba, int w, int h, std::string& err);
bool save_exr(const std::string& path, const float* rgba, int w, int h, std::string& err);
bool save_pfm(const std::string& path, const float* rgba, int w, int h, std::string& err);
bool save_png8(const std::string& path, const unsigned char* rgba8, int w, int h, std::string& err);
// by extension (.hdr / .exr / .pfm), as BitmapTexture::Save (texture.cpp:150-180); unknown extension -> .hdr rules
bool save_image(const std::string& path, const float* rgba, int w, int h, std::string& err);

}  // namespace pupil::host
