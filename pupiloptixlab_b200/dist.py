"""Multi-GPU rendering: sample-interleaved sharding + one reduce of the fp32 sum buffers.

One process per GPU (torchrun).  Every rank holds the whole scene and builds the same BVH on its own
device (the device build is deterministic, so replicas are identical and nothing is broadcast).
Frame (= sample index) s of pixel p is seeded Init(4, p, s) (example/path_tracer/main.cu:53,
pt_pass.cpp:45,56) no matter which GPU renders it, so rank r of N simply renders the frames
s = first + r, first + r + N, ...  — pb_render_params.sample_stride — and the per-rank SUM buffers
are added with ONE collective: `reduce(SUM)` to rank 0 over NCCL (NVLink 5 / NVSwitch).  There is no
other exchange on the path.  The reference has no multi-GPU code at all.

The same class runs over gloo with a CPU backend context in the tests (world_size 2, no GPU).
"""
from __future__ import annotations

import numpy as np

from . import capi


def shard_frames(first_sample: int, sample_count: int, rank: int, world: int) -> tuple[int, int, int]:
    """(first, count, stride) of the frames rank `rank` of `world` renders out of
    [first_sample, first_sample + sample_count): those with (s - first_sample) % world == rank."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    count = (sample_count - rank + world - 1) // world if sample_count > rank else 0
    return first_sample + rank, count, world


class ShardedRenderer:
    """Wraps one per-rank context (capi.Context on this rank's GPU) and the process group."""

    def __init__(self, ctx: capi.Context, rank: int | None = None, world: int | None = None, group=None):
        import torch.distributed as dist
        self.ctx, self.group = ctx, group
        self.dist = dist if dist.is_available() and dist.is_initialized() else None
        self.rank = rank if rank is not None else (self.dist.get_rank(group) if self.dist else 0)
        self.world = world if world is not None else (self.dist.get_world_size(group) if self.dist else 1)

    def render_device(self, width, height, max_depth, first_sample, sample_count, sum_tensor, stream=None, dst=0, flags=0):
        """Device path: `sum_tensor` is this rank's [h, w, 4] fp32 CUDA tensor; after the call rank `dst`
        holds the sum over all ranks (sum.xyz = radiance sums, sum.w = frame count)."""
        import torch
        first, count, stride = shard_frames(first_sample, sample_count, self.rank, self.world)
        p = capi.RenderParams(width, height, max_depth, first, count, stride, capi.RENDER_CLEAR | flags, 0)
        s = stream if stream is not None else torch.cuda.current_stream(sum_tensor.device)
        self.ctx.render_dev(p, sum_tensor.data_ptr(), 0, 0, s.cuda_stream)
        if self.world > 1:
            self.dist.reduce(sum_tensor, dst=dst, op=self.dist.ReduceOp.SUM, group=self.group)
        return sum_tensor

    def render_host(self, width, height, max_depth, first_sample, sample_count, dst=0):
        """Host-buffer path (numpy in, numpy out on rank dst; None elsewhere)."""
        import torch
        first, count, stride = shard_frames(first_sample, sample_count, self.rank, self.world)
        acc, _, _ = self.ctx.render(width, height, max_depth, first, count, stride, want_aov=False)
        if self.world > 1:
            t = torch.from_numpy(acc)
            if self.dist.get_backend(self.group) == "nccl":
                d = t.cuda()
                self.dist.reduce(d, dst=dst, op=self.dist.ReduceOp.SUM, group=self.group)
                t = d.cpu()
            else:
                self.dist.reduce(t, dst=dst, op=self.dist.ReduceOp.SUM, group=self.group)
            acc = t.numpy()
        return acc if self.rank == dst else None


def resolve(sum_rgba: np.ndarray) -> np.ndarray:
    """The reference's frame buffer: running mean of the frames (main.cu:187-193) = sum.xyz / sum.w, alpha 1."""
    out = np.ones_like(sum_rgba)
    w = sum_rgba[..., 3:4]
    out[..., :3] = np.where(w > 0, sum_rgba[..., :3] / np.maximum(w, 1), 0)
    return out
