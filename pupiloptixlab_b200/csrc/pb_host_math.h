// pb_host_math.h — host-side scene preparation arithmetic of the product (C++ host layer).
// world->object of an affine 3x4: adjugate / determinant evaluated in double, rounded once to
// float.  Stands in for DirectX::XMMatrixInverse as used by Mat4::GetInverse
// (framework/util/type.h:108-111) / the OptiX instance inverse; DirectXMath is not available here,
// so the expression is part of this framework's definition.
#pragma once
namespace pb {
inline void invert_affine(const float* M, float* inv) {
    const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], g = M[8], h = M[9], i = M[10];
    const double A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    const double det = a * A + b * B + c * C;
    const double r[9] = {A / det, -(b * i - c * h) / det, (b * f - c * e) / det,
                         B / det, (a * i - c * g) / det, -(a * f - c * d) / det,
                         C / det, -(a * h - b * g) / det, (a * e - b * d) / det};
    const double tx = M[3], ty = M[7], tz = M[11];
    for (int k = 0; k < 3; ++k) {
        inv[4 * k + 0] = (float)r[3 * k + 0];
        inv[4 * k + 1] = (float)r[3 * k + 1];
        inv[4 * k + 2] = (float)r[3 * k + 2];
        inv[4 * k + 3] = (float)(-(r[3 * k + 0] * tx + r[3 * k + 1] * ty + r[3 * k + 2] * tz));
    }
}
}  // namespace pb
