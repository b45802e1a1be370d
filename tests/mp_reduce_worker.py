"""Worker of tests/test_gpu_multi.py::test_two_ranks_over_nccl — one process per GPU under torch.distributed.run.
gloo only carries the 128-byte NCCL unique id and the per-rank films to rank 0; the reduce under test is the
ncclReduce that pb_render (PB_RENDER_REDUCE) enqueues itself.  Rank 0 writes the verdict as JSON to argv[1]."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pupiloptixlab_b200 import capi, scenes  # noqa: E402


def main():
    out_path = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dist.init_process_group("gloo")
    torch.cuda.set_device(local)
    api = capi.load()
    g = capi.Context(api, local)
    uid = [g.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    g.comm_init_rank(uid[0], world, rank)
    assert g.comm_info() == (world, rank)
    w, h, depth, S, first = 160, 120, 8, 3, 5
    scenes.cornell_box(g, w, h, "materials")
    # pre-reduce film of this rank: frames first + rank + k * world
    mine = g.render(w, h, depth, first + rank, S, world, want_aov=False)[0]
    # the collective under test, device buffers and host buffers
    dev = torch.device("cuda", local)
    d_sum = torch.zeros((h, w, 4), dtype=torch.float32, device=dev)
    p = capi.RenderParams(w, h, depth, first + rank, S, world, capi.RENDER_CLEAR | capi.RENDER_REDUCE, 0)
    g.render_dev(p, d_sum.data_ptr())
    torch.cuda.synchronize(dev)
    reduced = d_sum.cpu().numpy()
    h_sum = np.full((h, w, 4), 7.0, np.float32)
    p2 = capi.RenderParams(w, h, depth, first + rank, S, world, capi.RENDER_CLEAR | capi.RENDER_REDUCE | capi.RENDER_HOST_BUFFERS, 0)
    g.render_into(p2, h_sum)
    # accumulate a second step on top (no CLEAR): root adds, the others stay untouched
    p3 = capi.RenderParams(w, h, depth, first + world * S + rank, S, world, capi.RENDER_REDUCE, 0)
    g.render_dev(p3, d_sum.data_ptr())
    torch.cuda.synchronize(dev)
    accumulated = d_sum.cpu().numpy()
    films = [None] * world if rank == 0 else None
    dist.gather_object(mine, films, dst=0)
    untouched = [None] * world if rank == 0 else None
    dist.gather_object(bool((h_sum == 7.0).all()) if rank else True, untouched, dst=0)
    verdict = None
    if rank == 0:
        res = {"world": world, "nccl_ranks": g.comm_info()[0]}
        # the same shards rendered sequentially on THIS GPU: bit-identical to what the other GPUs produced
        seq = [g.render(w, h, depth, first + r, S, world, want_aov=False)[0] for r in range(world)]
        res["per_rank_bit_exact"] = [bool(np.array_equal(seq[r].view(np.uint32), films[r].view(np.uint32))) for r in range(world)]
        total = np.zeros_like(seq[0])
        for f in seq:
            total = total + f
        res["reduced_equals_sum_bits"] = bool(np.array_equal(reduced.view(np.uint32), total.view(np.uint32)))
        res["reduced_max_rel_vs_sum"] = float((np.abs(reduced - total) / (np.abs(total) + 1e-3)).max())
        res["host_buffers_equal_device"] = bool(np.array_equal(h_sum.view(np.uint32), reduced.view(np.uint32)))
        one = g.render(w, h, depth, first, S * world, 1, want_aov=False)[0]
        res["max_rel_vs_one_gpu"] = float((np.abs(reduced - one) / (np.abs(one) + 1e-3)).max())
        res["counts"] = [float(reduced[..., 3].min()), float(reduced[..., 3].max()), S * world]
        two = g.render(w, h, depth, first, 2 * S * world, 1, want_aov=False)[0]
        res["accumulated_max_rel_vs_one_gpu"] = float((np.abs(accumulated - two) / (np.abs(two) + 1e-3)).max())
        res["non_root_host_buffers_untouched"] = all(untouched)
        verdict = res
        json.dump(res, open(out_path, "w"))
    dist.barrier()
    g.close()
    dist.destroy_process_group()
    return 0 if verdict is None or all(verdict["per_rank_bit_exact"]) else 1


if __name__ == "__main__":
    sys.exit(main())
