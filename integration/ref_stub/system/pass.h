#pragma once
#include <cstdint>
#include <string>
#include <string_view>
namespace Pupil {
enum class EPassTag : uint32_t { None = 0, Pre = 1 << 0, Post = 1 << 1, Asyn = 1 << 2 };   // REF framework/system/pass.h:8 | None = 0,
class Pass {
public:
    const std::string name;                  // REF framework/system/pass.h:29 | const std::string name;
    const EPassTag tag;
    Pass(std::string_view name, EPassTag tag = EPassTag::None) noexcept : name(name), tag(tag) {}   // REF framework/system/pass.h:32 | Pass(std::string_view name, EPassTag tag = EPassTag::None) noexcept
    virtual ~Pass() = default;
    virtual void OnRun() noexcept = 0;       // REF framework/system/pass.h:38 | virtual void OnRun() noexcept = 0;
};
}  // namespace Pupil
