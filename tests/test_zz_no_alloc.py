"""-m gpu, collected LAST (file name): pb_trace_closest_static, the allocation-free closest-hit entry point added at the
request of the run's operators (no counterpart in the reference).  Its scratch is a file-scope device array
(pupiloptixlab_b200/csrc/pb_trace.cu), shared by every call and keeping the last call's contents, so the checks are:

* through the C ABI, twice in a row (a large batch, then a smaller different one: stale staging or a cursor that was not
  cleared would show), host buffers and device buffers: ids / t / u / v BIT-EXACT against pb_trace_closest — the
  tolerance every hit-id test of this repository uses;
* batches beyond PB_STATIC_TRACE_MAX_RAYS are refused with PB_ERR_LIMIT and a message that names the limit, the context
  stays usable; n == 0 is fine;
* no device memory is allocated by the calls (cudaMemGetInfo before / after).
Passes alone in a fresh process and after the rest of the suite (it builds its own context and scene)."""
import ctypes as C

import numpy as np
import pytest

from pupiloptixlab_b200 import capi, scenes
from tests import util_scenes as U

pytestmark = pytest.mark.gpu
MAX_RAYS = 262144          # PB_STATIC_TRACE_MAX_RAYS, include/pupil_b200.h


def static_closest(g, rays, host=True, dev_ptrs=None):
    hits = np.empty(len(rays), capi.HIT_DTYPE)
    if host:
        rc = g.api._trace_closest_static(g.h, rays.ctypes.data, len(rays), hits.ctypes.data, 1, None)
    else:
        rc = g.api._trace_closest_static(g.h, dev_ptrs[0], len(rays), dev_ptrs[1], 0, None)
    return rc, hits


def same(a, b):
    return all(np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)) for f in ("prim", "t", "u", "v"))


def test_static_entry_point_matches_pb_trace_closest_twice_in_a_row():
    import torch
    api = capi.load()
    g = capi.Context(api, 0)
    pos, idx, sph = U.random_soup(20000, 300, 17)
    U.oracle_scene(g, pos, idx, sph)             # same scene description helper the trace tests use (product context)
    first = np.ascontiguousarray(U.random_rays(200000, 31))
    first["tmax"][::3] = 1.5
    second = np.ascontiguousarray(U.random_rays(50001, 32, extent=1.0))
    want1, want2 = g.trace_closest(first), g.trace_closest(second)
    assert (want1["prim"] != capi.MISS).mean() > 0.2
    free0 = torch.cuda.mem_get_info(0)[0]
    # ---- host buffers, two calls in a row, then the first batch again
    for rays, want in ((first, want1), (second, want2), (first, want1)):
        rc, got = static_closest(g, rays)
        assert rc == 0, api._last_error(g.h)
        assert same(got, want)
    assert torch.cuda.mem_get_info(0)[0] == free0, "the static entry point must not allocate device memory"
    # ---- device buffers (allocated by the caller beforehand), two calls in a row
    for rays, want in ((second, want2), (first, want1)):
        d_r = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
        d_h = torch.zeros((len(rays), 4), dtype=torch.float32, device="cuda")
        rc = api._trace_closest_static(g.h, d_r.data_ptr(), len(rays), d_h.data_ptr(), 0, None)
        assert rc == 0, api._last_error(g.h)
        got = d_h.cpu().numpy().view(capi.HIT_DTYPE).reshape(-1)
        assert same(got, want)
    # ---- the limit
    big = np.ascontiguousarray(U.random_rays(MAX_RAYS + 1, 33))
    rc, _ = static_closest(g, big)
    assert rc == 4                                             # PB_ERR_LIMIT
    msg = api._last_error(g.h).decode()
    assert "PB_STATIC_TRACE_MAX_RAYS" in msg and str(MAX_RAYS) in msg
    rc, got = static_closest(g, big[:MAX_RAYS])                 # exactly the maximum is served
    assert rc == 0 and same(got, g.trace_closest(big[:MAX_RAYS]))
    rc, _ = static_closest(g, big[:0])
    assert rc == 0
    # ---- a second context on the same device shares the array: sequential calls still agree
    g2 = capi.Context(api, 0)
    scenes.cornell_box(g2, 32, 32, "materials")
    r3 = np.ascontiguousarray(U.random_rays(30000, 34, extent=0.9))
    r3["oy"] += 1.0
    rc, got = static_closest(g2, r3)
    assert rc == 0 and same(got, g2.trace_closest(r3))
    rc, got = static_closest(g, second)
    assert rc == 0 and same(got, want2)


def test_static_entry_point_needs_a_committed_scene():
    g = capi.Context(capi.load(), 0)
    rays = np.ascontiguousarray(U.random_rays(16, 1))
    rc, _ = static_closest(g, rays)
    assert rc == 5                                             # PB_ERR_NOT_COMMITTED
