// scripts/chaos_nudge.h — force-included (-include) when scripts/chaos_floor.py compiles a second copy of the TEST ORACLE:
// sinf / cosf results move by one ulp for a pseudo-random half of their arguments, which is what separates libdevice from glibc.
#include <cmath>
#include <math.h>
#include <cstring>
#include <cstdint>
// emulate libdevice-vs-glibc: results of sinf / cosf off by one ulp for a pseudo-random half of the arguments
static inline float orc_nudge(float r, float x) {
    uint32_t b; std::memcpy(&b, &x, 4);
    b ^= b >> 15; b *= 0x2c1b3c6du; b ^= b >> 12;
    if (b & 1u) return r;
    return (b & 2u) ? nextafterf(r, 2.f) : nextafterf(r, -2.f);
}
static inline float nudged_sinf(float x) { return orc_nudge(::sinf(x), x); }
static inline float nudged_cosf(float x) { return orc_nudge(::cosf(x), x); }
#define sinf nudged_sinf
#define cosf nudged_cosf
