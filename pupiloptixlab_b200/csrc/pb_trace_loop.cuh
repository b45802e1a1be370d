// pb_trace_loop.cuh — the persistent-warp ray loop shared by the ray-cast kernels (pb_trace.cu) and
// the wavefront extend / connect kernels (pb_wavefront.cu).
//
// One ray per lane, but lanes do NOT wait for the slowest ray of their batch: whenever fewer than
// `refill` lanes of the warp still hold a ray, the idle lanes claim new rays from the global cursor
// (one atomicAdd per refill, warp-aggregated) and traversal resumes with a full warp ("dynamic
// fetch", Aila & Laine 2009).  Without it the measured lane utilisation of the node loop was 7.7 / 32
// on the 10 M-triangle scene (profiles/r01_extend_v0.md).
//
// IO supplies the per-ray plumbing:
//   bool load(uint32_t i, V3& o, V3& d, float& tmin, float& tmax)   ray i of the queue (false = skip)
//   void store(uint32_t i, const Hit& h)                             result of ray i (any lane, any time)
//   void flush()                                                     warp-converged point after a batch of stores
//   static constexpr bool kFixedRange                                every ray runs over [kPathTmin, kPathTmax] (path rays)
#pragma once
#include "pb_traverse.cuh"

namespace pb {

// PB_DEFER_STORE (default): a finished ray's result stays in the lane's registers until the warp's next refill and is
// written there, together with the other finished lanes' (>= 32 - refill of them), instead of in the iteration it
// finished in.  ncu on round 1's kernel: the finish path (record store + the material lookup that bins the path) ran
// in 22.7 M of 29.9 M warp iterations at 2.4 active lanes; the lane is idle until the refill anyway.
#ifndef PB_DEFER_STORE
#define PB_DEFER_STORE 1
#endif

template<bool ANY, bool COUNT, class IO>
__device__ __forceinline__ void trace_loop(const TraceScene& sc, IO& io, LaneStack& st, uint32_t* cursor, uint32_t n,
                                           uint32_t refill, TraceCounters& tc) {
    const unsigned lane = threadIdx.x & 31u;
    Traversal<ANY, COUNT, IO::kFixedRange> tr;
    bool have = false, exhausted = false, pending = false;   // pending: tr.best of ray idx is finished and not yet stored
    uint32_t idx = 0;
    while (true) {
        if (!exhausted) {
            const unsigned need = __ballot_sync(0xFFFFFFFFu, !have);
            if (need) {
                const int leader = __ffs(need) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)__popc(need));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (!have) {
                    if (PB_DEFER_STORE && pending) { io.store(idx, tr.result(st)); pending = false; }
                    idx = base + (uint32_t)__popc(need & ((1u << lane) - 1u));
                    if (idx < n) {
                        V3 o, d; float tmin, tmax;
                        if (io.load(idx, o, d, tmin, tmax)) {
                            tr.begin(sc, o, d, tmin, tmax, st);
                            if (tr.finished_at_begin()) io.store(idx, tr.result(st));
                            else have = true;
                        }
                    }
                }
                if (base + (uint32_t)__popc(need) >= n) exhausted = true;   // warp-uniform
                if (PB_DEFER_STORE) io.flush();    // bins what was just stored, before a lane can finish (and store) again
            }
        }
        if (!__any_sync(0xFFFFFFFFu, have)) {
            if (PB_DEFER_STORE && pending) { io.store(idx, tr.result(st)); pending = false; }
            io.flush();
            if (exhausted) break;
            continue;
        }
        while (true) {
            if (tr.step(sc, st, &tc, have)) {   // all lanes call step()
                have = false;
                if (PB_DEFER_STORE) pending = true; else io.store(idx, tr.result(st));
            }
            const unsigned m = __ballot_sync(0xFFFFFFFFu, have);
            if (m == 0u || (!exhausted && (uint32_t)__popc(m) < refill)) break;
        }
        if (!PB_DEFER_STORE) io.flush();
        else if (exhausted) {   // no refill will come: write what finished, bin it
            if (pending) { io.store(idx, tr.result(st)); pending = false; }
            io.flush();
        }
    }
}

}  // namespace pb
