#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <string_view>
#include <unordered_map>
#include "util/util.h"
typedef unsigned long long CUdeviceptr;      // <cuda.h>
namespace Pupil {
enum class EBufferFlag : unsigned int {
    None = 0,                                // REF framework/system/buffer.h:16 | None = 0,
    SharedWithDX12 = 1,
    AllowDisplay = 1 << 1,                   // REF framework/system/buffer.h:18 | AllowDisplay = 1 << 1,
};
struct BufferDesc {
    const char* name = nullptr;              // REF framework/system/buffer.h:27 | const char *name = nullptr;
    EBufferFlag flag = EBufferFlag::None;    // REF framework/system/buffer.h:28 | EBufferFlag flag = EBufferFlag::None;
    uint32_t width = 1;                      // REF framework/system/buffer.h:29 | uint32_t width = 1;
    uint32_t height = 1;                     // REF framework/system/buffer.h:30 | uint32_t height = 1;
    uint32_t stride_in_byte = 1;             // REF framework/system/buffer.h:31 | uint32_t stride_in_byte = 1;
};
struct Buffer {
    BufferDesc desc{};
    CUdeviceptr cuda_ptr = 0;                // REF framework/system/buffer.h:36 | CUdeviceptr cuda_ptr = 0;
};
class BufferManager : public util::Singleton<BufferManager> {
public:
    constexpr static std::string_view DEFAULT_FINAL_RESULT_BUFFER_NAME = "final result";   // REF framework/system/buffer.h:46 | constexpr static std::string_view DEFAULT_FINAL_RESULT_BUFFER_NAME = "final result";
    [[nodiscard]] Buffer* GetBuffer(std::string_view) noexcept;     // REF framework/system/buffer.h:53 | [[nodiscard]] Buffer *GetBuffer(std::string_view) noexcept;
    Buffer* AllocBuffer(const BufferDesc&) noexcept;                // REF framework/system/buffer.h:54 | Buffer *AllocBuffer(const BufferDesc &) noexcept;
private:
    std::unordered_map<std::string, std::unique_ptr<Buffer>> m_buffers;
};
}  // namespace Pupil
