// image_io.cpp — see image_io.h.  Part 1: byte-level readers (PNG, BMP, TGA, PNM), the LDR decode of the
// reference, Radiance HDR and PFM.  Part 2 (below): OpenEXR subset, writers.
#include "image_io.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace pupil::host {
namespace {

// Untrusted headers: every decoder bounds its dimensions before it multiplies or allocates (no size_t overflow, no
// multi-gigabyte allocation for a 30-byte file).  2^15 x 2^15 texels covers any texture or environment map.
bool dims_ok(long long w, long long h) { return w > 0 && h > 0 && w <= 32768 && h <= 32768; }


struct Bytes {            // decoded 8-bit image, `c` interleaved channels (1 grey, 2 grey+alpha, 3 rgb, 4 rgba)
    int w = 0, h = 0, c = 0;
    std::vector<unsigned char> px;
};

uint32_t be32(const unsigned char* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
uint32_t le32(const unsigned char* p) { return (uint32_t)p[3] << 24 | (uint32_t)p[2] << 16 | (uint32_t)p[1] << 8 | p[0]; }
uint32_t le16(const unsigned char* p) { return (uint32_t)p[1] << 8 | p[0]; }

bool zlib_inflate(const unsigned char* src, size_t n, std::vector<unsigned char>& dst, size_t expect) {
    dst.resize(expect);
    uLongf len = (uLongf)expect;
    int rc = uncompress(dst.data(), &len, src, (uLong)n);
    if (rc != Z_OK) return false;
    dst.resize(len);
    return true;
}

// ---- PNG (ISO/IEC 15948): all colour types, bit depths 1-16, Adam7; 16-bit samples keep their high byte -----------
int paeth(int a, int b, int c) {
    int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
bool png_unfilter(unsigned char* data, size_t rows, size_t stride, size_t bpp, std::string& err) {
    // data = rows x (1 + stride); reconstructs in place, previous row starts as zeros
    std::vector<unsigned char> zero(stride, 0);
    const unsigned char* prev = zero.data();
    for (size_t y = 0; y < rows; ++y) {
        unsigned char* row = data + y * (stride + 1);
        const int ft = row[0];
        unsigned char* cur = row + 1;
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= bpp ? cur[x - bpp] : 0, b = prev[x], c = x >= bpp ? prev[x - bpp] : 0;
            int v = cur[x];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: err = "png: bad filter type"; return false;
            }
            cur[x] = (unsigned char)v;
        }
        prev = cur;
    }
    return true;
}
bool read_png(const unsigned char* d, size_t n, Bytes& out, std::string& err) {
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 || memcmp(d, sig, 8)) { err = "png: bad signature"; return false; }
    size_t pos = 8;
    uint32_t W = 0, H = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    while (pos + 12 <= n) {
        const uint32_t len = be32(d + pos);
        const unsigned char* tag = d + pos + 4;
        const unsigned char* body = d + pos + 8;
        if (pos + 12 + (size_t)len > n) { err = "png: truncated chunk"; return false; }
        if (!memcmp(tag, "IHDR", 4) && len >= 13) {
            W = be32(body); H = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
        } else if (!memcmp(tag, "PLTE", 4)) plte.assign(body, body + len);
        else if (!memcmp(tag, "tRNS", 4)) trns.assign(body, body + len);
        else if (!memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!memcmp(tag, "IEND", 4)) break;
        pos += 12 + (size_t)len;
    }
    if (!W || !H || W > 32768 || H > 32768) { err = "png: bad size"; return false; }
    const int chans = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!chans || !(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) { err = "png: unsupported colour type / depth"; return false; }
    const size_t bits_pp = (size_t)chans * depth, bpp = std::max<size_t>(1, bits_pp / 8);
    auto stride_of = [&](size_t w) { return (w * bits_pp + 7) / 8; };
    // pass geometry (Adam7) or the single full pass
    struct Pass { int x0, y0, dx, dy; };
    static const Pass adam7[7] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
    const Pass full = {0, 0, 1, 1};
    const int n_pass = interlace ? 7 : 1;
    size_t raw = 0;
    for (int p = 0; p < n_pass; ++p) {
        const Pass& ps = interlace ? adam7[p] : full;
        const size_t pw = (W - ps.x0 + ps.dx - 1) / ps.dx, ph = (H - ps.y0 + ps.dy - 1) / ps.dy;
        if (pw && ph) raw += ph * (1 + stride_of(pw));
    }
    std::vector<unsigned char> buf;
    if (!zlib_inflate(idat.data(), idat.size(), buf, raw) || buf.size() != raw) { err = "png: inflate failed"; return false; }
    // samples -> 8-bit `chans` per pixel (palette indices kept for now)
    std::vector<unsigned char> s8((size_t)W * H * chans);
    size_t off = 0;
    const int scale = depth == 1 ? 255 : depth == 2 ? 85 : depth == 4 ? 17 : 1;
    for (int p = 0; p < n_pass; ++p) {
        const Pass& ps = interlace ? adam7[p] : full;
        if ((int)W <= ps.x0 || (int)H <= ps.y0) continue;
        const size_t pw = (W - ps.x0 + ps.dx - 1) / ps.dx, ph = (H - ps.y0 + ps.dy - 1) / ps.dy, st = stride_of(pw);
        if (!png_unfilter(buf.data() + off, ph, st, bpp, err)) return false;
        for (size_t y = 0; y < ph; ++y) {
            const unsigned char* row = buf.data() + off + y * (st + 1) + 1;
            for (size_t x = 0; x < pw; ++x) {
                unsigned char* o = &s8[(((size_t)ps.y0 + y * ps.dy) * W + ps.x0 + x * ps.dx) * chans];
                for (int k = 0; k < chans; ++k) {
                    const size_t si = x * chans + k;
                    if (depth == 8) o[k] = row[si];
                    else if (depth == 16) o[k] = row[2 * si];           // high byte
                    else {
                        const size_t bit = si * depth;
                        const int v = (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                        o[k] = (unsigned char)(ctype == 3 ? v : v * scale);
                    }
                }
            }
        }
        off += ph * (st + 1);
    }
    out.w = (int)W; out.h = (int)H;
    if (ctype == 3) {          // palette -> rgb / rgba
        const bool alpha = !trns.empty();
        out.c = alpha ? 4 : 3;
        out.px.resize((size_t)W * H * out.c);
        for (size_t i = 0; i < (size_t)W * H; ++i) {
            const size_t idx = s8[i];
            for (int k = 0; k < 3; ++k) out.px[i * out.c + k] = 3 * idx + k < plte.size() ? plte[3 * idx + k] : 0;
            if (alpha) out.px[i * 4 + 3] = idx < trns.size() ? trns[idx] : 255;
        }
    } else {
        out.c = chans;
        out.px.swap(s8);
    }
    return true;
}

// ---- BMP (BITMAPINFOHEADER family, uncompressed 8 / 24 / 32 bit) ---------------------------------------------------

bool read_bmp(const unsigned char* d, size_t n, Bytes& out, std::string& err) {
    if (n < 54 || d[0] != 'B' || d[1] != 'M') { err = "bmp: bad header"; return false; }
    const uint32_t data_off = le32(d + 10), hdr = le32(d + 14);
    const int32_t W = (int32_t)le32(d + 18), Hs = (int32_t)le32(d + 22);
    const uint32_t bpp = le16(d + 28), comp = le32(d + 30);
    if (hdr < 40 || Hs == INT32_MIN || !(comp == 0 || comp == 3) || !(bpp == 8 || bpp == 24 || bpp == 32)) { err = "bmp: unsupported variant"; return false; }
    const int H = Hs < 0 ? -Hs : Hs;
    if (!dims_ok(W, H)) { err = "bmp: bad dimensions"; return false; }
    const size_t stride = ((size_t)W * bpp / 8 + 3) & ~(size_t)3;
    if ((size_t)data_off > n || stride * (size_t)H > n - data_off) { err = "bmp: truncated"; return false; }
    // 8-bit files index a 256-entry BGRA palette right behind the info header: all of it must be inside the file
    if (bpp == 8 && ((size_t)hdr > n || 14 + (size_t)hdr + 4 * 256 > n)) { err = "bmp: palette outside the file"; return false; }
    const unsigned char* pal = d + 14 + hdr;
    out.w = W; out.h = H; out.c = bpp == 32 ? 4 : 3;
    out.px.resize((size_t)W * H * out.c);
    for (int y = 0; y < H; ++y) {
        const unsigned char* row = d + data_off + stride * (size_t)(Hs < 0 ? y : H - 1 - y);
        for (int x = 0; x < W; ++x) {
            unsigned char* o = &out.px[((size_t)y * W + x) * out.c];
            if (bpp == 8) { const unsigned char* e = pal + 4 * row[x]; o[0] = e[2]; o[1] = e[1]; o[2] = e[0]; }
            else { const unsigned char* s = row + (size_t)x * (bpp / 8); o[0] = s[2]; o[1] = s[1]; o[2] = s[0]; if (bpp == 32) o[3] = s[3]; }
        }
    }
    return true;
}

// ---- TGA (types 2 / 3 true-colour & grey, 10 / 11 run-length) -------------------------------------------------------
bool read_tga(const unsigned char* d, size_t n, Bytes& out, std::string& err) {
    if (n < 18) { err = "tga: truncated"; return false; }
    const int id_len = d[0], cmap = d[1], type = d[2], W = (int)le16(d + 12), H = (int)le16(d + 14), bpp = d[16], desc = d[17];
    if (cmap || !(type == 2 || type == 3 || type == 10 || type == 11) || !dims_ok(W, H) || !(bpp == 8 || bpp == 24 || bpp == 32)) { err = "tga: unsupported variant"; return false; }
    const int bytes = bpp / 8;
    out.w = W; out.h = H; out.c = bytes == 1 ? 1 : bytes;
    out.px.resize((size_t)W * H * out.c);
    size_t pos = 18 + (size_t)id_len;
    std::vector<unsigned char> raw((size_t)W * H * bytes);
    if (type == 2 || type == 3) {
        if (pos + raw.size() > n) { err = "tga: truncated"; return false; }
        memcpy(raw.data(), d + pos, raw.size());
    } else {
        size_t o = 0;
        while (o < raw.size()) {
            if (pos >= n) { err = "tga: truncated"; return false; }
            const int hd = d[pos++], cnt = (hd & 127) + 1;
            if (hd & 128) {
                if (pos + bytes > n) { err = "tga: truncated"; return false; }
                for (int k = 0; k < cnt && o < raw.size(); ++k, o += bytes) memcpy(&raw[o], d + pos, bytes);
                pos += bytes;
            } else {
                const size_t len = std::min((size_t)cnt * bytes, raw.size() - o);
                if (pos + len > n) { err = "tga: truncated"; return false; }
                memcpy(&raw[o], d + pos, len);
                pos += len; o += len;
            }
        }
    }
    const bool top = (desc & 0x20) != 0;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const unsigned char* s = &raw[((size_t)(top ? y : H - 1 - y) * W + x) * bytes];
            unsigned char* o = &out.px[((size_t)y * W + x) * out.c];
            if (bytes == 1) o[0] = s[0];
            else { o[0] = s[2]; o[1] = s[1]; o[2] = s[0]; if (bytes == 4) o[3] = s[3]; }
        }
    return true;
}

// ---- PNM (P5 grey / P6 colour, maxval <= 255) ------------------------------------------------------------------------
bool pnm_token(const unsigned char* d, size_t n, size_t& pos, long& v) {
    while (pos < n) {
        if (d[pos] == '#') { while (pos < n && d[pos] != '\n') ++pos; }
        else if (isspace(d[pos])) ++pos;
        else break;
    }
    if (pos >= n || !isdigit(d[pos])) return false;
    v = 0;
    while (pos < n && isdigit(d[pos])) { v = v * 10 + (d[pos++] - '0'); if (v > 100000000L) return false; }   // no signed overflow on digit runs
    return true;
}
bool read_pnm(const unsigned char* d, size_t n, Bytes& out, std::string& err) {
    const int c = d[1] == '5' ? 1 : 3;
    size_t pos = 2;
    long W, H, mx;
    if (!pnm_token(d, n, pos, W) || !pnm_token(d, n, pos, H) || !pnm_token(d, n, pos, mx) || mx <= 0 || mx > 255 || !dims_ok(W, H)) { err = "pnm: bad header"; return false; }
    ++pos;   // the single whitespace after maxval
    if (pos > n || (size_t)W * H * c > n - pos) { err = "pnm: truncated"; return false; }
    out.w = (int)W; out.h = (int)H; out.c = c;
    out.px.assign(d + pos, d + pos + (size_t)W * H * c);
    if (mx != 255) for (auto& v : out.px) v = (unsigned char)((int)v * 255 / mx);
    return true;
}

// StbImageLoad's 8-bit branch, texture.cpp:107-118: `c` channels as in the file, channels 0..2 read at i, i+1, i+2
// whatever c is (so grey images pick up the neighbouring pixels - reproduced, clamped at the end of the buffer)
void ldr_to_float(const Bytes& b, Image& out) {
    out.w = b.w; out.h = b.h;
    out.rgba.resize((size_t)b.w * b.h * 4);
    const size_t data_size = (size_t)b.w * b.h * b.c, last = data_size - 1;
    size_t j = 0;
    for (size_t i = 0; i < data_size; i += b.c) {
        for (int k = 0; k < 3; ++k) out.rgba[j++] = (float)pow(b.px[std::min(i + k, last)] * 1.f / 255.f, 2.2f);
        out.rgba[j++] = b.c == 4 ? b.px[i + 3] * 1.f / 255.f : 1.f;
    }
}

// ---- Radiance .hdr (RGBE, flat or new-style run-length scanlines) ----------------------------------------------------
bool read_hdr(const unsigned char* d, size_t n, Image& out, std::string& err) {
    size_t pos = 0;
    auto line = [&](std::string& s) { s.clear(); while (pos < n && d[pos] != '\n') s.push_back((char)d[pos++]); if (pos < n) ++pos; return pos <= n; };
    std::string l;
    line(l);
    if (l.rfind("#?RADIANCE", 0) != 0 && l.rfind("#?RGBE", 0) != 0) { err = "hdr: bad signature"; return false; }
    bool fmt = false;
    while (pos < n) { line(l); if (l.empty()) break; if (l.find("FORMAT=32-bit_rle_rgbe") != std::string::npos) fmt = true; }
    if (!fmt) { err = "hdr: unsupported format"; return false; }
    line(l);
    int W = 0, H = 0;
    if (sscanf(l.c_str(), "-Y %d +X %d", &H, &W) != 2 || !dims_ok(W, H)) { err = "hdr: unsupported orientation"; return false; }
    out.w = W; out.h = H;
    out.rgba.resize((size_t)W * H * 4);
    std::vector<unsigned char> scan((size_t)W * 4);
    for (int y = 0; y < H; ++y) {
        bool rle = false;
        if (W >= 8 && W < 32768 && pos + 4 <= n && d[pos] == 2 && d[pos + 1] == 2 && !(d[pos + 2] & 0x80) && ((d[pos + 2] << 8) | d[pos + 3]) == W) rle = true;
        if (rle) {
            pos += 4;
            for (int k = 0; k < 4; ++k) {
                int x = 0;
                while (x < W) {
                    if (pos >= n) { err = "hdr: truncated"; return false; }
                    int cnt = d[pos++];
                    if (cnt > 128) {
                        cnt -= 128;
                        if (pos >= n || x + cnt > W) { err = "hdr: bad run"; return false; }
                        const unsigned char v = d[pos++];
                        for (int i = 0; i < cnt; ++i) scan[(size_t)(x++) * 4 + k] = v;
                    } else {
                        if (cnt == 0 || pos + cnt > n || x + cnt > W) { err = "hdr: bad run"; return false; }
                        for (int i = 0; i < cnt; ++i) scan[(size_t)(x++) * 4 + k] = d[pos++];
                    }
                }
            }
        } else {
            if (pos + (size_t)W * 4 > n) { err = "hdr: truncated"; return false; }
            memcpy(scan.data(), d + pos, (size_t)W * 4);
            pos += (size_t)W * 4;
        }
        for (int x = 0; x < W; ++x) {
            const unsigned char* p = &scan[(size_t)x * 4];
            float* o = &out.rgba[((size_t)y * W + x) * 4];
            if (p[3]) { const float f = ldexpf(1.0f, (int)p[3] - (128 + 8)); o[0] = p[0] * f; o[1] = p[1] * f; o[2] = p[2] * f; }
            else o[0] = o[1] = o[2] = 0.f;
            o[3] = 1.f;   // c == 3 for .hdr: alpha 1 (texture.cpp:100-104)
        }
    }
    return true;
}

// ---- PFM ("PF" colour / "Pf" grey, little or big endian, bottom row first) -------------------------------------------
bool read_pfm(const unsigned char* d, size_t n, Image& out, std::string& err) {
    if (n < 3) { err = "pfm: truncated"; return false; }
    const int c = d[1] == 'F' ? 3 : 1;
    int W = 0, H = 0, used = 0;
    float scale = 0.f;
    // the header is text inside a binary file that is not NUL-terminated: parse a bounded, terminated copy of it
    char head[96];
    const size_t hn = std::min(n - 2, sizeof head - 1);
    memcpy(head, d + 2, hn);
    head[hn] = 0;
    if (sscanf(head, "%d %d %f%n", &W, &H, &scale, &used) != 3 || !dims_ok(W, H)) { err = "pfm: bad header"; return false; }
    size_t pos = 2 + (size_t)used + 1;
    if (pos > n || (size_t)W * H * c * 4 > n - pos) { err = "pfm: truncated"; return false; }
    const bool little = scale < 0.f;
    out.w = W; out.h = H;
    out.rgba.resize((size_t)W * H * 4);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float v[3];
            for (int k = 0; k < c; ++k) {
                const unsigned char* p = d + pos + (((size_t)(H - 1 - y) * W + x) * c + k) * 4;
                uint32_t u = little ? le32(p) : be32(p);
                memcpy(&v[k], &u, 4);
            }
            float* o = &out.rgba[((size_t)y * W + x) * 4];
            o[0] = v[0]; o[1] = c == 3 ? v[1] : v[0]; o[2] = c == 3 ? v[2] : v[0]; o[3] = 1.f;
        }
    return true;
}

bool read_exr(const unsigned char* d, size_t n, Image& out, std::string& err);   // part 2

// ---- OpenEXR subset: single-part scanline files, compression NONE / ZIPS / ZIP, HALF / FLOAT / UINT channels named
// R G B A (or one luminance channel).  LoadEXR of tinyexr returns RGBA float, row 0 at the top (texture.cpp:130-147).
float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 31, man = h & 1023, u;
    if (exp == 0) {
        if (man == 0) u = sign;
        else { int e = -1; do { ++e; man <<= 1; } while (!(man & 1024)); u = sign | (uint32_t)(127 - 15 - e) << 23 | (man & 1023) << 13; }
    } else if (exp == 31) u = sign | 0x7f800000u | man << 13;
    else u = sign | (exp + 112) << 23 | man << 13;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
bool read_exr(const unsigned char* d, size_t n, Image& out, std::string& err) {
    if (n < 8 || le32(d) != 20000630u) { err = "exr: bad magic"; return false; }
    const uint32_t ver = le32(d + 4);
    if ((ver & 0xff) != 2 || (ver & 0x1a00)) { err = "exr: tiled / multi-part / deep files are not supported"; return false; }
    size_t pos = 8;
    struct Chan { std::string name; int type; };
    std::vector<Chan> chans;
    int comp = -1, x0 = 0, y0 = 0, x1 = -1, y1 = -1;
    auto cstr = [&](std::string& s) { s.clear(); while (pos < n && d[pos]) s.push_back((char)d[pos++]); ++pos; return pos <= n; };
    while (pos < n) {
        std::string name, type;
        cstr(name);
        if (name.empty()) break;
        cstr(type);
        if (pos + 4 > n) { err = "exr: truncated header"; return false; }
        const uint32_t size = le32(d + pos);
        pos += 4;
        if ((size_t)size > n - pos) { err = "exr: truncated header"; return false; }
        const unsigned char* v = d + pos;
        if (name == "channels") {
            size_t q = 0;
            while (q < size && v[q]) {
                Chan c;
                while (q < size && v[q]) c.name.push_back((char)v[q++]);
                ++q;
                if (q + 16 > size) { err = "exr: bad channel list"; return false; }
                c.type = (int)le32(v + q);
                if (le32(v + q + 8) != 1 || le32(v + q + 12) != 1) { err = "exr: subsampled channels are not supported"; return false; }
                q += 16;
                chans.push_back(c);
            }
        } else if (name == "compression" && size >= 1) comp = v[0];
        else if (name == "dataWindow" && size >= 16) { x0 = (int)le32(v); y0 = (int)le32(v + 4); x1 = (int)le32(v + 8); y1 = (int)le32(v + 12); }
        pos += size;
    }
    if (chans.empty() || x1 < x0 || y1 < y0) { err = "exr: missing channels / dataWindow"; return false; }
    if (!(comp == 0 || comp == 2 || comp == 3)) { err = "exr: only NONE / ZIPS / ZIP compression is supported"; return false; }
    if (!dims_ok((long long)x1 - x0 + 1, (long long)y1 - y0 + 1) || chans.size() > 64) { err = "exr: bad dataWindow / channel count"; return false; }
    const int W = x1 - x0 + 1, H = y1 - y0 + 1, lines = comp == 3 ? 16 : 1, n_chunks = (H + lines - 1) / lines;
    if (pos > n || (size_t)n_chunks * 8 > n - pos) { err = "exr: truncated offset table"; return false; }
    size_t line_bytes = 0;
    std::vector<size_t> ch_off(chans.size());
    for (size_t c = 0; c < chans.size(); ++c) { ch_off[c] = line_bytes; line_bytes += (size_t)W * (chans[c].type == 1 ? 2 : 4); }
    int slot[4] = {-1, -1, -1, -1};   // R G B A -> channel index
    for (size_t c = 0; c < chans.size(); ++c) {
        const std::string& nm = chans[c].name;
        const char last = nm.empty() ? 0 : nm.back();
        const bool plain = nm.size() == 1 || (nm.size() > 1 && nm[nm.size() - 2] == '.');
        if (!plain) continue;
        if (last == 'R') slot[0] = (int)c; else if (last == 'G') slot[1] = (int)c; else if (last == 'B') slot[2] = (int)c; else if (last == 'A') slot[3] = (int)c;
        else if (last == 'Y' && slot[0] < 0) slot[0] = slot[1] = slot[2] = (int)c;
    }
    if (slot[0] < 0 && chans.size() == 1) slot[0] = slot[1] = slot[2] = 0;
    if (slot[0] < 0 || slot[1] < 0 || slot[2] < 0) { err = "exr: no R, G, B channels"; return false; }
    out.w = W; out.h = H;
    out.rgba.assign((size_t)W * H * 4, 1.f);
    std::vector<unsigned char> raw, tmp;
    for (int k = 0; k < n_chunks; ++k) {
        uint64_t off = (uint64_t)le32(d + pos + 8 * (size_t)k) | (uint64_t)le32(d + pos + 8 * (size_t)k + 4) << 32;
        if (off > n || n - off < 8) { err = "exr: bad chunk offset"; return false; }
        const long long yl = (long long)(int)le32(d + off) - y0;
        const uint32_t sz = le32(d + off + 4);
        if ((uint64_t)sz > n - off - 8 || yl < 0 || yl >= H) { err = "exr: bad chunk"; return false; }
        const int y = (int)yl;
        const int rows = std::min(lines, H - y);
        const size_t want = line_bytes * rows;
        const unsigned char* src = d + off + 8;
        if (comp != 0 && sz < want) {
            if (!zlib_inflate(src, sz, tmp, want) || tmp.size() != want) { err = "exr: inflate failed"; return false; }
            for (size_t i = 1; i < want; ++i) tmp[i] = (unsigned char)(tmp[i - 1] + tmp[i] - 128);      // predictor
            raw.resize(want);
            const size_t half = (want + 1) / 2;
            for (size_t i = 0; i < want; ++i) raw[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];          // de-interleave
            src = raw.data();
        } else if (sz != want) { err = "exr: unexpected chunk size"; return false; }
        for (int r = 0; r < rows; ++r)
            for (int s = 0; s < 4; ++s) {
                if (slot[s] < 0) continue;
                const Chan& c = chans[slot[s]];
                const unsigned char* p = src + line_bytes * r + ch_off[slot[s]];
                for (int x = 0; x < W; ++x) {
                    float v;
                    if (c.type == 1) v = half_to_float((uint16_t)le16(p + 2 * x));
                    else if (c.type == 2) { uint32_t u = le32(p + 4 * x); memcpy(&v, &u, 4); }
                    else v = (float)le32(p + 4 * x);
                    out.rgba[((size_t)(y + r) * W + x) * 4 + s] = v;
                }
            }
    }
    return true;
}

void put32(std::vector<unsigned char>& b, uint32_t v) { for (int k = 0; k < 4; ++k) b.push_back((unsigned char)(v >> (8 * k))); }
void put_str(std::vector<unsigned char>& b, const char* s) { while (*s) b.push_back((unsigned char)*s++); b.push_back(0); }
void put_f32(std::vector<unsigned char>& b, float f) { uint32_t u; memcpy(&u, &f, 4); put32(b, u); }
bool write_file(const std::string& path, const std::vector<unsigned char>& b, std::string& err) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { err = "cannot open [" + path + "] for writing"; return false; }
    const bool ok = fwrite(b.data(), 1, b.size(), f) == b.size();
    fclose(f);
    if (!ok) err = "short write to [" + path + "]";
    return ok;
}

bool ends_with(const std::string& s, const char* suf) {
    const size_t m = strlen(suf);
    if (s.size() < m) return false;
    for (size_t i = 0; i < m; ++i) if (tolower((unsigned char)s[s.size() - m + i]) != suf[i]) return false;
    return true;
}

}  // namespace

bool load_image_memory(const unsigned char* d, size_t n, const std::string& name, Image& out, std::string& err) {
    if (n < 4) { err = "image '" + name + "': file too short"; return false; }
    Bytes b;
    bool ok = false, ldr = false;
    if (ends_with(name, ".exr") || (d[0] == 0x76 && d[1] == 0x2f && d[2] == 0x31 && d[3] == 0x01)) ok = read_exr(d, n, out, err);
    else if (d[0] == 0x89 && d[1] == 'P') { ok = read_png(d, n, b, err); ldr = true; }
    else if (d[0] == '#' && d[1] == '?') ok = read_hdr(d, n, out, err);
    else if (d[0] == 'P' && (d[1] == 'F' || d[1] == 'f')) ok = read_pfm(d, n, out, err);
    else if (d[0] == 'P' && (d[1] == '5' || d[1] == '6')) { ok = read_pnm(d, n, b, err); ldr = true; }
    else if (d[0] == 'B' && d[1] == 'M') { ok = read_bmp(d, n, b, err); ldr = true; }
    else if (d[0] == 0xff && d[1] == 0xd8) { err = "JPEG decoding is not built in (convert the texture to PNG)"; ok = false; }
    else if (ends_with(name, ".tga")) { ok = read_tga(d, n, b, err); ldr = true; }
    else err = "unknown image format";
    if (!ok) { err = "image '" + name + "': " + err; return false; }
    if (ldr) ldr_to_float(b, out);
    return true;
}

bool load_image(const std::string& path, Image& out, std::string& err) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { err = "fail to load image [" + path + "]"; return false; }
    std::vector<unsigned char> buf;
    unsigned char tmp[1 << 16];
    size_t k;
    while ((k = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + k);
    fclose(f);
    return load_image_memory(buf.data(), buf.size(), path, out, err);
}

// SaveExr, texture.cpp:23-86: three 32-bit float channels B, G, R; image rows written bottom-up (y = h-1 first)
bool save_exr(const std::string& path, const float* rgba, int w, int h, std::string& err) {
    std::vector<unsigned char> b;
    put32(b, 20000630u); put32(b, 2u);
    put_str(b, "channels"); put_str(b, "chlist"); put32(b, 3 * 18 + 1);
    for (const char* nm : {"B", "G", "R"}) { put_str(b, nm); put32(b, 2u); put32(b, 0u); put32(b, 1u); put32(b, 1u); }
    b.push_back(0);
    put_str(b, "compression"); put_str(b, "compression"); put32(b, 1); b.push_back(0);
    for (const char* nm : {"dataWindow", "displayWindow"}) { put_str(b, nm); put_str(b, "box2i"); put32(b, 16); put32(b, 0); put32(b, 0); put32(b, (uint32_t)(w - 1)); put32(b, (uint32_t)(h - 1)); }
    put_str(b, "lineOrder"); put_str(b, "lineOrder"); put32(b, 1); b.push_back(0);
    put_str(b, "pixelAspectRatio"); put_str(b, "float"); put32(b, 4); put_f32(b, 1.f);
    put_str(b, "screenWindowCenter"); put_str(b, "v2f"); put32(b, 8); put_f32(b, 0.f); put_f32(b, 0.f);
    put_str(b, "screenWindowWidth"); put_str(b, "float"); put32(b, 4); put_f32(b, 1.f);
    b.push_back(0);
    const size_t line = (size_t)w * 12, table = b.size();
    for (int y = 0; y < h; ++y) { const uint64_t off = table + (size_t)h * 8 + (size_t)y * (8 + line); put32(b, (uint32_t)off); put32(b, (uint32_t)(off >> 32)); }
    for (int y = 0; y < h; ++y) {
        put32(b, (uint32_t)y); put32(b, (uint32_t)line);
        const float* row = rgba + (size_t)(h - 1 - y) * w * 4;
        for (int c : {2, 1, 0}) for (int x = 0; x < w; ++x) put_f32(b, row[4 * x + c]);
    }
    return write_file(path, b, err);
}

// SaveHdr, texture.cpp:13-22: stbi_write_hdr of the first three channels with stbi_flip_vertically_on_write(true);
// RGBE conversion as stb_image_write does it (frexp of the largest component), flat (non run-length) scanlines
bool save_hdr(const std::string& path, const float* rgba, int w, int h, std::string& err) {
    std::vector<unsigned char> b;
    char hdr[128];
    const int m = snprintf(hdr, sizeof hdr, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    b.insert(b.end(), hdr, hdr + m);
    for (int y = h - 1; y >= 0; --y)
        for (int x = 0; x < w; ++x) {
            const float* p = rgba + ((size_t)y * w + x) * 4;
            const float mx = std::max(p[0], std::max(p[1], p[2]));
            if (!(mx >= 1e-32f)) { for (int k = 0; k < 4; ++k) b.push_back(0); continue; }
            int e;
            const float norm = (float)frexp(mx, &e) * 256.0f / mx;
            b.push_back((unsigned char)(p[0] * norm)); b.push_back((unsigned char)(p[1] * norm)); b.push_back((unsigned char)(p[2] * norm));
            b.push_back((unsigned char)(e + 128));
        }
    return write_file(path, b, err);
}

bool save_pfm(const std::string& path, const float* rgba, int w, int h, std::string& err) {
    std::vector<unsigned char> b;
    char hdr[64];
    const int m = snprintf(hdr, sizeof hdr, "PF\n%d %d\n-1.0\n", w, h);
    b.insert(b.end(), hdr, hdr + m);
    for (int y = h - 1; y >= 0; --y)
        for (int x = 0; x < w; ++x)
            for (int k = 0; k < 3; ++k) put_f32(b, rgba[((size_t)y * w + x) * 4 + k]);
    return write_file(path, b, err);
}

bool save_png8(const std::string& path, const unsigned char* rgba8, int w, int h, std::string& err) {
    std::vector<unsigned char> raw((size_t)h * (1 + (size_t)w * 4));
    for (int y = 0; y < h; ++y) {
        raw[(size_t)y * (1 + (size_t)w * 4)] = 0;
        memcpy(&raw[(size_t)y * (1 + (size_t)w * 4) + 1], rgba8 + (size_t)y * w * 4, (size_t)w * 4);
    }
    uLongf zn = compressBound((uLong)raw.size());
    std::vector<unsigned char> z(zn);
    if (compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6) != Z_OK) { err = "png: deflate failed"; return false; }
    z.resize(zn);
    std::vector<unsigned char> b = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    auto be = [&](uint32_t v) { for (int k = 3; k >= 0; --k) b.push_back((unsigned char)(v >> (8 * k))); };
    auto chunk = [&](const char* tag, const std::vector<unsigned char>& body) {
        be((uint32_t)body.size());
        const size_t at = b.size();
        b.insert(b.end(), tag, tag + 4);
        b.insert(b.end(), body.begin(), body.end());
        be((uint32_t)crc32(0L, b.data() + at, (uInt)(b.size() - at)));
    };
    std::vector<unsigned char> ihdr;
    for (uint32_t v : {(uint32_t)w, (uint32_t)h}) for (int k = 3; k >= 0; --k) ihdr.push_back((unsigned char)(v >> (8 * k)));
    for (unsigned char v : {8, 6, 0, 0, 0}) ihdr.push_back(v);
    chunk("IHDR", ihdr); chunk("IDAT", z); chunk("IEND", {});
    return write_file(path, b, err);
}

bool save_image(const std::string& path, const float* rgba, int w, int h, std::string& err) {
    if (w <= 0 || h <= 0 || !rgba) { err = "empty image"; return false; }
    if (ends_with(path, ".exr")) return save_exr(path, rgba, w, h, err);
    if (ends_with(path, ".pfm")) return save_pfm(path, rgba, w, h, err);
    return save_hdr(path, rgba, w, h, err);
}

}  // namespace pupil::host
