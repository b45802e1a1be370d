// pb_slots.cuh — where a path of a wave lives: wave slot <-> (frame k, path j) <-> pixel.  PB_HD so that tests/hostsim can
// check the bijections on the CPU for any film size.
#pragma once
#include "pb_types.h"

namespace pb {

// Path j of a frame -> pixel.  A frame is walked in strips of PB_TILE_ROWS rows, column by column, so that 32 consecutive
// paths (one warp of every stage, one refill of the traversal loop) cover an 8 x 4 pixel block instead of a 32 x 1 row:
// their primary rays share more nodes and their bounce rays start closer together.  A bijection for any film size (the
// last strip just has fewer rows); seeds and sums are per pixel, so results do not depend on it.
#ifndef PB_TILE_ROWS
#define PB_TILE_ROWS 4u
#endif
PB_HD uint32_t path_pixel(uint32_t j, uint32_t w, uint32_t h) {
    if (PB_TILE_ROWS <= 1u) return j;
    const uint32_t strips = h / PB_TILE_ROWS, full = strips * PB_TILE_ROWS * w;
    if (j < full) {
        const uint32_t s = j / (PB_TILE_ROWS * w), q = j - s * PB_TILE_ROWS * w;
        return (PB_TILE_ROWS * s + q % PB_TILE_ROWS) * w + q / PB_TILE_ROWS;
    }
    const uint32_t r = h - strips * PB_TILE_ROWS, q = j - full;
    return (strips * PB_TILE_ROWS + q % r) * w + q / r;
}

// Slot of a wave <-> (frame k of the wave, path j of the frame).  The frames of a wave are interleaved in groups of
// PB_FRAME_GROUP paths: slots [g * G * K, (g + 1) * G * K) hold paths [g * G, (g + 1) * G) of all K frames, frame by frame,
// so the rays in flight at any moment come from a K times smaller part of the film than with frame-major slots and
// their leaves and primitives stay in L2.  Any film size (the last group is shorter); PB_FRAME_GROUP 0 = frame-major.
#ifndef PB_FRAME_GROUP
#define PB_FRAME_GROUP 256u   // C4, same box: frame-major 492.9, groups of 32 / 256 / 4096 paths 494.5 / 495.2 / 495.1 Msamples/s
#endif
PB_HD void slot_to_kj(uint32_t slot, uint32_t n_px, uint32_t K, uint32_t& k, uint32_t& j) {
    if (PB_FRAME_GROUP == 0u) { k = slot / n_px; j = slot - k * n_px; return; }
    const uint32_t G = PB_FRAME_GROUP, full = n_px / G, body = full * G * K;
    if (slot < body) { const uint32_t g = slot / (G * K), rem = slot - g * G * K; k = rem / G; j = g * G + (rem - k * G); return; }
    const uint32_t r = n_px - full * G, rem = slot - body;
    k = rem / r; j = full * G + (rem - k * r);
}
PB_HD uint32_t kj_to_slot(uint32_t k, uint32_t j, uint32_t n_px, uint32_t K) {
    if (PB_FRAME_GROUP == 0u) return k * n_px + j;
    const uint32_t G = PB_FRAME_GROUP, full = n_px / G, g = j / G;
    if (g < full) return g * G * K + k * G + (j - g * G);
    return full * G * K + k * (n_px - full * G) + (j - full * G);
}

}  // namespace pb
