"""Multi-GPU host logic on CPU: world_size-2 gloo, the per-rank renderer is the test oracle (the
product needs a GPU).  Sample-interleaved shards + one reduce(SUM) must reproduce the single-rank
render: frame counts exactly, radiance sums up to the float re-association of (a+c)+(b+d) vs
((a+b)+c)+d, and every frame is rendered exactly once."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pupiloptixlab_b200 import dist as pdist


def test_shard_frames_partitions_every_frame_once():
    for first in (0, 5):
        for count in (0, 1, 2, 7, 64):
            for world in (1, 2, 3, 4, 8):
                seen = []
                for r in range(world):
                    f, c, s = pdist.shard_frames(first, count, r, world)
                    assert s == world
                    seen += [f + i * s for i in range(c)]
                assert sorted(seen) == list(range(first, first + count))
    with pytest.raises(ValueError):
        pdist.shard_frames(0, 4, 2, 2)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle
    from pupiloptixlab_b200 import scenes
    pyoracle.load()._set_threads(2)
    o = pyoracle.OracleContext(pyoracle.load())
    scenes.cornell_box(o, 48, 40, "materials")
    r = pdist.ShardedRenderer(o)
    assert (r.rank, r.world) == (rank, world)
    acc = r.render_host(48, 40, 6, 3, 5)          # frames 3..7 over 2 ranks: 3 + 2
    if rank == 0:
        np.save(out_path, acc)
    else:
        assert acc is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_render_equals_single_rank(tmp_path, oracle_api):
    out = str(tmp_path / "acc.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    from oracle import pyoracle
    from pupiloptixlab_b200 import scenes
    o = pyoracle.OracleContext(oracle_api)
    scenes.cornell_box(o, 48, 40, "materials")
    want, _, _ = o.render(48, 40, 6, 3, 5, want_aov=False)
    assert np.array_equal(got[..., 3], want[..., 3]) and (got[..., 3] == 5).all()
    np.testing.assert_allclose(got[..., :3], want[..., :3], rtol=2e-6, atol=1e-6)
    # the shards themselves are bit-identical to single frames rendered alone (determinism)
    a, _, _ = o.render(48, 40, 6, 3, 3, 2, want_aov=False)   # rank 0: frames 3, 5, 7
    b, _, _ = o.render(48, 40, 6, 4, 2, 2, want_aov=False)   # rank 1: frames 4, 6
    assert np.array_equal((a + b).view(np.uint32), got.view(np.uint32))
    img = pdist.resolve(got)
    assert img.shape == got.shape and (img[..., 3] == 1).all()
