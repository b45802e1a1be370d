// ORACLE shim (test infrastructure): util::Singleton as used by framework/cuda/texture.h.
#pragma once
#include <memory>
#include <vector>
namespace Pupil::util {
template<typename T>
class Singleton {
public:
    static T* instance() { static const std::unique_ptr<T> i = std::make_unique<T>(); return i.get(); }
protected:
    Singleton() {}
};
}  // namespace Pupil::util
