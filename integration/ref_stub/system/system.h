#pragma once
namespace Pupil {
enum class ESystemEvent {
    Quit, Precompute, StartRendering, StopRendering,
    SceneLoad,                               // REF framework/system/system.h:18 | SceneLoad,
    FrameFinished
};
}  // namespace Pupil
