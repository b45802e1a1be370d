#pragma once
#include <memory>
namespace Pupil::util {
template<typename T>
class Singleton {
public:
    static T* instance() {          // REF framework/util/util.h:13 | static T *instance() {
        static const std::unique_ptr<T> instance = std::make_unique<T>();
        return instance.get();
    }
protected:
    Singleton() {}
};
}  // namespace Pupil::util
