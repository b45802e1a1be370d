// ORACLE shim (test infrastructure).  The reference's framework/cuda/random.h is included VERBATIM
// (via #include_next) under the name RandomAsWritten; Pupil::cuda::Random derives from it and only
// re-states Next2() with the argument evaluation order the reference has where it actually runs:
//   reference:  return make_float2(Next(), Next());          (random.h:38-40)
// C++ leaves the order of the two Next() calls unspecified.  nvcc device code (the only place the
// reference runs) evaluates left to right — x gets the first draw (verified: `nvcc -ptx` of
// Next2() stores {first, second}) — while g++ on the host evaluates right to left.
// Init()/Next()/GetSeed()/SetSeed() are the reference's own code, untouched.
#pragma once
#define Random RandomAsWritten
#include_next "cuda/random.h"
#undef Random
namespace Pupil::cuda {
class Random : public RandomAsWritten {
public:
    float2 Next2() noexcept {
        float a = Next();
        float b = Next();
        return make_float2(a, b);
    }
};
}  // namespace Pupil::cuda
