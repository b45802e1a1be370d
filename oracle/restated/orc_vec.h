// ORACLE — test infrastructure, not product code.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
//
// orc_vec.h — minimal fp32 vector algebra with the SAME operation order as the reference's
// framework/cuda/vec_math.h (dot :462, cross :467, length :472, normalize :477, v/s as
// v*(1/s) :425-428, lerp :439) so that results are bit-identical to the reference headers when
// both are compiled for the host without FMA contraction.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

constexpr float kPi = 3.14159265358979323846f;     // M_PIf   vec_math.h:43
constexpr float kInvPi = 0.318309886183790671538f; // M_1_PIf vec_math.h:49

struct f2 { float x, y; };
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };

inline f3 mk3(float s) { return {s, s, s}; }
inline f3 mk3(float x, float y, float z) { return {x, y, z}; }
inline f2 mk2(float x, float y) { return {x, y}; }

inline f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline f3 operator-(f3 a) { return {-a.x, -a.y, -a.z}; }
inline f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline f3 operator*(f3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline f3 operator*(float s, f3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline f3 operator/(f3 a, f3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline f3 operator/(f3 a, float s) { float inv = 1.0f / s; return a * inv; } // vec_math.h:425
inline f3 operator-(float s, f3 a) { return {s - a.x, s - a.y, s - a.z}; }
inline f3& operator+=(f3& a, f3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
inline f3& operator*=(f3& a, f3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
inline f3& operator*=(f3& a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
inline f3& operator/=(f3& a, float s) { float inv = 1.0f / s; a *= inv; return a; } // :432

inline f2 operator+(f2 a, f2 b) { return {a.x + b.x, a.y + b.y}; }
inline f2 operator*(f2 a, float s) { return {a.x * s, a.y * s}; }
inline f2 operator*(float s, f2 a) { return {a.x * s, a.y * s}; }

inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(f4 a, f4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline f3 cross(f3 a, f3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline float length(f3 v) { return sqrtf(dot(v, v)); }
inline f3 normalize(f3 v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }
inline f4 normalize(f4 v) { float inv = 1.0f / sqrtf(dot(v, v)); return {v.x * inv, v.y * inv, v.z * inv, v.w * inv}; }
inline f3 lerp(f3 a, f3 b, float t) { return a + t * (b - a); }

inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

}  // namespace orc
