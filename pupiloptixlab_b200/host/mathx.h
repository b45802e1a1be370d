// mathx.h — host-side matrix / transform / camera math of the C++ host layer.
//
// Mirrors Pupil::util::{Mat4, Transform, Camera} (framework/util/type.h:73-113,
// framework/util/transform.cpp:7-98, framework/util/camera.cpp:7-101).  The reference implements
// these on DirectXMath (XMMatrixPerspectiveFovRH, XMMatrixLookAtRH, XMMatrixInverse,
// XMMatrixTranspose), which is a Windows SDK header that is not vendored; they are restated here
// from their public definitions, evaluated in double and rounded once to float ("parity unpinned",
// DESIGN.md §1).  Mat4 is row-major with the column-vector convention: translation in e[3], e[7],
// e[11] (transform.cpp:100-106).
#pragma once
#include <cmath>
#include <cstring>

namespace pupil::host {

struct Float3 { float x = 0, y = 0, z = 0; };

struct Mat4 {
    float e[16];
    Mat4() { std::memset(e, 0, sizeof(e)); }
    static Mat4 identity() { Mat4 m; m.e[0] = m.e[5] = m.e[10] = m.e[15] = 1.f; return m; }
    float& at(int r, int c) { return e[4 * r + c]; }
    float at(int r, int c) const { return e[4 * r + c]; }
};

// Mat4::operator* (type.h:97): plain matrix product
inline Mat4 mul(const Mat4& a, const Mat4& b) {
    Mat4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += (double)a.at(i, k) * (double)b.at(k, j);
            r.at(i, j) = (float)s;
        }
    return r;
}
inline Mat4 transpose(const Mat4& a) {
    Mat4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.at(i, j) = a.at(j, i);
    return r;
}
// general 4x4 inverse (Gauss-Jordan with partial pivoting, double) — XMMatrixInverse stand-in
inline bool inverse_d(const double* m, double* out) {
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { a[i][j] = m[4 * i + j]; a[i][4 + j] = i == j ? 1.0 : 0.0; }
    for (int c = 0; c < 4; ++c) {
        int p = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
        if (a[p][c] == 0.0) return false;
        if (p != c) for (int j = 0; j < 8; ++j) { double t = a[c][j]; a[c][j] = a[p][j]; a[p][j] = t; }
        const double inv = 1.0 / a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] *= inv;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            const double f = a[r][c];
            if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[4 * i + j] = a[i][4 + j];
    return true;
}
inline Mat4 inverse(const Mat4& m) {
    double d[16], o[16];
    for (int i = 0; i < 16; ++i) d[i] = m.e[i];
    Mat4 r = Mat4::identity();
    if (inverse_d(d, o)) for (int i = 0; i < 16; ++i) r.e[i] = (float)o[i];
    return r;
}

// util::Transform: every operation LEFT-multiplies (transform.cpp:49,67,84)
struct Transform {
    Mat4 matrix = Mat4::identity();

    void Rotate(float ux, float uy, float uz, float angle) {  // quaternion form, transform.cpp:7-50
        double n = std::sqrt((double)ux * ux + (double)uy * uy + (double)uz * uz);
        double x = ux / n, y = uy / n, z = uz / n;
        double th = (double)angle / 180.0 * 3.14159265358979323846;
        double a = std::cos(0.5 * th), s = std::sin(0.5 * th);
        double b = s * x, c = s * y, d = s * z;
        Mat4 r = Mat4::identity();
        r.at(0, 0) = (float)(1 - 2 * c * c - 2 * d * d); r.at(0, 1) = (float)(2 * b * c - 2 * a * d); r.at(0, 2) = (float)(2 * a * c + 2 * b * d);
        r.at(1, 0) = (float)(2 * b * c + 2 * a * d); r.at(1, 1) = (float)(1 - 2 * b * b - 2 * d * d); r.at(1, 2) = (float)(2 * c * d - 2 * a * b);
        r.at(2, 0) = (float)(2 * b * d - 2 * a * c); r.at(2, 1) = (float)(2 * a * b + 2 * c * d); r.at(2, 2) = (float)(1 - 2 * b * b - 2 * c * c);
        matrix = mul(r, matrix);
    }
    void Translate(float x, float y, float z) {
        Mat4 t = Mat4::identity();
        t.at(0, 3) = x; t.at(1, 3) = y; t.at(2, 3) = z;
        matrix = mul(t, matrix);
    }
    void Scale(float x, float y, float z) {
        Mat4 s = Mat4::identity();
        s.at(0, 0) = x; s.at(1, 1) = y; s.at(2, 2) = z;
        matrix = mul(s, matrix);
    }
    // XMMatrixLookAtRH then transpose(inverse(.)) (transform.cpp:86-98): camera-to-world whose
    // columns are the camera axes (x = up × z, z = eye - target) and the eye position
    void LookAt(Float3 o, Float3 t, Float3 up) {
        double zx = (double)o.x - t.x, zy = (double)o.y - t.y, zz = (double)o.z - t.z;
        double n = std::sqrt(zx * zx + zy * zy + zz * zz); zx /= n; zy /= n; zz /= n;
        double xx = (double)up.y * zz - (double)up.z * zy, xy = (double)up.z * zx - (double)up.x * zz, xz = (double)up.x * zy - (double)up.y * zx;
        n = std::sqrt(xx * xx + xy * xy + xz * xz); xx /= n; xy /= n; xz /= n;
        double yx = zy * xz - zz * xy, yy = zz * xx - zx * xz, yz = zx * xy - zy * xx;
        Mat4 m = Mat4::identity();
        m.at(0, 0) = (float)xx; m.at(1, 0) = (float)xy; m.at(2, 0) = (float)xz;
        m.at(0, 1) = (float)yx; m.at(1, 1) = (float)yy; m.at(2, 1) = (float)yz;
        m.at(0, 2) = (float)zx; m.at(1, 2) = (float)zy; m.at(2, 2) = (float)zz;
        m.at(0, 3) = o.x; m.at(1, 3) = o.y; m.at(2, 3) = o.z;
        matrix = m;
    }
    // "Mitsuba 3: +X left, +Z view; Pupil: +X right, +Z -view" — negate columns 0 and 2 of the
    // upper 3 rows (util_loader.cpp:161-166, scene.cpp:134-139)
    void FlipXZ() { for (int r = 0; r < 3; ++r) { matrix.at(r, 0) *= -1.f; matrix.at(r, 2) *= -1.f; } }
};

// util::Camera (camera.cpp): the two matrices the device camera needs (render/camera.h:7-10)
struct Camera {
    float fov_y = 90.f, aspect = 1.f, near_clip = 0.01f, far_clip = 10000.f;
    Transform to_world;

    // transpose(inverse(PerspectiveFovRH * Translation(1,1,0) * Scaling(.5,.5,1))), camera.cpp:7-20
    Mat4 GetSampleToCameraMatrix() const {
        const float fov = fov_y / 180.f * 3.14159265358979323846f;   // float, as the reference passes it
        const double h = 1.0 / std::tan((double)fov * 0.5), w = h / (double)aspect;
        const double zn = near_clip, zf = far_clip, r = zf / (zn - zf);
        const double proj[16] = {w, 0, 0, 0, 0, h, 0, 0, 0, 0, r, -1, 0, 0, r * zn, 0};
        const double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 1, 1, 0, 1};
        const double S[16] = {0.5, 0, 0, 0, 0, 0.5, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        double pt[16], m[16], inv[16];
        auto mm = [](const double* a, const double* b, double* o) {
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += a[4 * i + k] * b[4 * k + j]; o[4 * i + j] = s; }
        };
        mm(proj, T, pt); mm(pt, S, m);
        Mat4 out = Mat4::identity();
        if (inverse_d(m, inv)) for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out.at(i, j) = (float)inv[4 * j + i];
        return out;
    }
    // SetWorldTransform keeps the matrix it was given (camera.cpp:80-101: m_to_world = to_world,
    // m_to_world_dirty = false), so GetToWorldMatrix returns it unchanged until the camera moves
    Mat4 GetToWorldMatrix() const { return to_world.matrix; }
};

}  // namespace pupil::host
