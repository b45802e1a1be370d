#pragma once
#include <functional>
#include <vector>
namespace Pupil {
template<auto enum_v>
struct StubEvent { static std::vector<std::function<void(void*)>>& ops() { static std::vector<std::function<void(void*)>> v; return v; } };
template<auto enum_v, typename Func>
inline void EventBinder(Func&& f) {    // REF framework/util/event.h:36 | inline void EventBinder(Func &&f) {
    StubEvent<enum_v>::ops().push_back(std::forward<Func>(f));
}
template<auto enum_v>
inline void EventDispatcher(void* p) { // REF framework/util/event.h:45 | inline void EventDispatcher(void *p) {
    for (auto& op : StubEvent<enum_v>::ops()) op(p);
}
}  // namespace Pupil
