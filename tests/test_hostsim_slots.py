"""Wave slot <-> (frame, path) <-> pixel (pupiloptixlab_b200/csrc/pb_slots.cuh) on the CPU through tests/hostsim: both
maps must be bijections for ANY film size and wave length, and k_accumulate's inverse must agree with k_generate's map."""
import ctypes as C

import numpy as np
import pytest


@pytest.mark.parametrize("w,h,K", [(1, 1, 1), (1, 7, 3), (7, 1, 2), (3, 3, 5), (37, 23, 3), (64, 48, 2), (256, 4, 1),
                                   (257, 5, 4), (16, 16, 32), (100, 9, 7), (33, 31, 1), (1920, 4, 2)])
def test_slot_maps_are_bijections(hostsim_lib, w, h, K):
    n_px = w * h
    frame = np.empty(K * n_px, np.uint32); pixel = np.empty(K * n_px, np.uint32); inv = np.empty(K * n_px, np.uint32)
    hostsim_lib.hs_slot_tables.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    hostsim_lib.hs_slot_tables(w, h, K, frame.ctypes.data, pixel.ctypes.data, inv.ctypes.data)
    assert frame.max() == K - 1 and pixel.max() == n_px - 1
    # every (frame, pixel) pair exactly once
    key = frame.astype(np.int64) * n_px + pixel
    assert np.array_equal(np.sort(key), np.arange(K * n_px))
    # the inverse map lands on the slot whose (frame, path) it was asked for: pixel_of(slot_of(k, j)) == path_pixel(j)
    slots = inv.reshape(K, n_px)
    assert np.array_equal(np.sort(inv), np.arange(K * n_px))
    assert (frame[slots] == np.arange(K, dtype=np.uint32)[:, None]).all()
    assert (pixel[slots] == pixel[slots[0]][None, :]).all()          # same path j -> same pixel in every frame
    if w % 8 == 0 and h % 4 == 0 and n_px % 256 == 0:               # a warp's 32 consecutive slots = one 8 x 4 block of one frame
        p = pixel[:32].astype(np.int64)
        assert len(set(frame[:32])) == 1 and (p % w).max() - (p % w).min() == 7 and (p // w).max() - (p // w).min() == 3
