#!/usr/bin/env python
"""Writes INTEGRATION.md; §1 quotes integration/b200_pt_pass.{h,cpp} verbatim (tests/test_integration_shim.py checks it)."""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rd = lambda *p: open(os.path.join(ROOT, *p)).read().strip()

MD = f"""# INTEGRATION — plugging libpupil_b200.so into PupilOptixLab

The reference's extension point for this path is a C++ subclass of `Pupil::Pass`
(`framework/system/pass.h:22-43`) registered with `System::AddPass` (`framework/system/system.cpp:129-134`)
and fed by the `SceneLoad` event (`system.cpp:164`, `example/path_tracer/pt_pass.cpp:211-223`).
`example/cuda_test/pass.cpp:13-78` is the reference's own precedent for a non-OptiX CUDA pass.
`libpupil_b200.so` is a C ABI (`include/pupil_b200.h`), so the binding a maintainer adds is one
translation unit that includes the header and links the library — no FFI generator is involved
(the reference is C++; an `extern "C"` header *is* its FFI).

## 1. The reference-side binding (what a maintainer adds to `example/path_tracer`)

The two files below are `integration/b200_pt_pass.h` and `integration/b200_pt_pass.cpp`, quoted verbatim.  They are
written against the reference's real types — `world->scene->shape_instances` (`framework/resource/scene.h:46`,
`resource::ShapeInstance` `framework/resource/shape.h:62-74`: `shape`, `mat`, `is_emitter`, `emitter`, `transform`,
`flip_*`), `resource::Material`'s tagged union (`framework/resource/material.h:66-84`), `scene->emitters`
(`scene.h:47`), `BufferDesc{{name, flag, width, height, stride_in_byte}}` and `Buffer::cuda_ptr`
(`framework/system/buffer.h:25-40`), `world->camera->GetSampleToCameraMatrix()` (`framework/world/camera.h:41-43`) —
and they are **compiled and run** by `tests/test_integration_shim.py`:

* `-fsyntax-only -Wall -Wextra -Werror` against `integration/ref_stub/`, headers that declare exactly the reference
  members used, each with a `// REF file:line | text` note that the same test checks against `/root/reference` where
  that tree exists (110+ notes), so the stubs — and with them this document — cannot drift from the reference again;
* linked against a fake `World` / `BufferManager` and a recording stand-in for the library: the test asserts the call
  sequence (instances in `World::LoadScene` order, unusable shapes skipped as `world.cpp:122` does, flags, material
  fields per type, environment emitter, camera, the four output buffers with the reference's names / strides / flags,
  `pb_update_instance` with the moved `RenderObject`'s index, frame and seed bookkeeping of `pt_pass.cpp:39-57`).

`main.cpp` (`example/path_tracer/main.cpp:5-21`) then does
`auto pt_pass = std::make_unique<Pupil::pt::B200PTPass>("Path Tracing"); system->AddPass(pt_pass.get());` instead of
`PTPass`; nothing else in the reference changes.  The OptiX module / pipeline / SBT objects (`framework/optix/*`) are
simply not instantiated for this pass.  Link: `-lpupil_b200 -lcudart`.

```cpp
{rd("integration", "b200_pt_pass.h")}
```

```cpp
{rd("integration", "b200_pt_pass.cpp")}
```

Asynchronous use (the reference's `optixLaunch` is asynchronous too, `framework/optix/pass.h:79-88`): with
`PB_RENDER_ASYNC` in `flags`, `pb_render_ex` returns once the work is enqueued on the stream; `Pass::Synchronize`
becomes `pb_wait(m_ctx)` (and `pb_get_stats` implies it).  Several renders can be queued back to back — with
host buffers too (pinned memory; uploads and downloads are enqueued on the same stream) — and their
ray counts and stage times are folded into one `pb_stats`.  Each wave of a render (generate, `max_depth` ×
{{extend, miss, 7 × shade, connect}}, accumulate — ~195 kernel launches) is captured once as a CUDA graph and
replayed with one `cudaGraphLaunch` per call (`PB_NO_GRAPH=1` disables this; the graph is keyed on the film size,
depth, sample stride, output pointers and a scene epoch that `pb_scene_commit` / `pb_set_camera` bump; a transform
refit by `pb_update_instance` keeps every device pointer and therefore every captured graph).

## 2. The headless stand-in shipped here

Because the reference cannot be built on Linux (MSVC/DX12/OptiX, SURVEY §8c), this repository ships the
same binding against a headless mirror of the interfaces it touches, in C++ under
`pupiloptixlab_b200/host/` (built by `__graft_entry__.build()` into `libpupil_host.so`):

| Reference interface | Mirror (same names / argument meaning / error behaviour) |
|---|---|
| `Pupil::Pass`, `System::{{AddPass,SetScene,Run}}` headless branch (`system.cpp:55-59`) | `host/system.h` |
| `BufferManager::{{AllocBuffer,GetBuffer}}` (`framework/system/buffer.h:34-59`), cudaMalloc branch only | `host/system.h` |
| `world::World` (scene, camera, `GetRenderobjects`), `world::RenderObject::{{UpdateTransform,ApplyTransform}}` raising `RenderInstanceTransform` → `RenderInstanceUpdate` (`render_object.cpp:46-54`, `world.cpp:45-58`) | `host/system.h` |
| `resource::Scene::LoadFromXML` + xml parser (`framework/resource/scene.cpp:27-224`, `xml/*`) | `host/scene_xml.{{h,cpp}}` |
| `util::Camera`, `util::Transform` (DirectXMath semantics restated) | `host/mathx.h` |
| `BitmapTexture::{{Load,Save}}` (stb / tinyexr in the reference) | `host/image_io.{{h,cpp}}` |
| `B200PTPass` (§1), on one or several GPUs | `host/b200_pt_pass.{{h,cpp}}` |

C entry points for tools and tests: `include/pupil_host.h` (`ph_scene_*`, `ph_run_path_tracer`, `ph_session_*`).
Python (`pupiloptixlab_b200/capi.py`, `hostlib.py`) is a `ctypes` view of the two ABIs for tests and benches:

```python
from pupiloptixlab_b200 import capi, scenes
ctx = capi.Context(capi.load(), device=0)     # raises without a GPU: no CPU fallback
scenes.cornell_box(ctx, 1024, 1024, "materials")
sum_rgba, albedo, normal = ctx.render(1024, 1024, max_depth=8, first_sample=0, sample_count=64)
```

## 3. Multi-GPU

The reference is single-GPU (`framework/cuda/context.cpp:15-44`).  Here the scene and BVH are replicated per GPU, frames
are dealt by sample index (frame `f` of pixel `p` is seeded `Init(4, p, f)` wherever it runs) and ONE
`ncclReduce(sum, fp32, root 0)` of the accumulation film per `pb_render` is issued **by the library**, on the render's
own stream (`pupiloptixlab_b200/csrc/pb_comm.cu`; NCCL is bound with `dlopen`, so a process that already carries a
libnccl — torch — keeps a single copy):

* one process per GPU (`bench.py --gpus N` under `torchrun`, or MPI): rank 0 calls `pb_comm_unique_id`, the 128 bytes
  travel by any means, every rank calls `pb_comm_init_rank(ctx, id, n, rank)` and then
  `pb_render(…, first_sample = base + rank, sample_stride = n, flags |= PB_RENDER_REDUCE)`; only the root's accumulation
  buffer is written;
* one process, n GPUs (what a C++ host of the reference does): `pb_create_multi(ctxs, devices, n)` +
  `pb_render_multi(ctxs, n, params_of_the_whole_job, targets_on_ctx0, streams)`; `B200PTPass(system, {{0, 1, …}})` and
  `ph_session_create_multi` sit on top of it (`tests/cpp/test_multi_gpu.cpp` is a plain C++ program using only the two
  C headers);
* callers that cannot call `cudaMalloc` (serving loops, harnesses that forbid it) have `pb_trace_closest_static`: the same
  closest-hit result with the kernel's scratch in a static device array — at most `PB_STATIC_TRACE_MAX_RAYS` = 262 144
  rays per call (12.0 MiB per GPU, held until exit), `PB_ERR_LIMIT` beyond, calls must not overlap on the GPU.  It has no
  counterpart in the reference and was added at the request of the run's operators;
* asynchronous NCCL errors are polled in `pb_wait` / `pb_get_stats` / synchronous renders and surface as `PB_ERR_NCCL`
  with NCCL's message in `pb_last_error`.
"""

open(os.path.join(ROOT, "INTEGRATION.md"), "w").write(MD)
print("wrote INTEGRATION.md", len(MD.splitlines()), "lines")
