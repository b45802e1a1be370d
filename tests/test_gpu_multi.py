"""-m gpu: the multi-GPU path on hardware.  Skips where the box has a single GPU, except for the 1-rank
communicator case, which still walks pb_create_multi / ncclReduce / the root's film add.

* tests/cpp/test_multi_gpu (C++, no Python, no torch): per-rank films bit-identical across GPUs, pb_render_multi ==
  their sum == the 1-GPU film up to float re-association, accumulation, PB_RENDER_ASYNC, B200PTPass on n devices;
* two processes under torch.distributed.run, one per GPU, NCCL communicator built by pb_comm_init_rank from a unique
  id shipped over gloo: pre-reduce buffers == the same shards rendered sequentially on one GPU BIT FOR BIT, reduced
  film == their sum, == the 1-GPU film within float re-association."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_multi_gpu")
XML = os.path.join(ROOT, "tests", "data", "cornell_materials.xml")


def n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("n", [1, 2, 4])
def test_cpp_host_on_n_gpus(n):
    assert os.path.exists(BIN), "tests/cpp/test_multi_gpu is missing: run __graft_entry__.build()"
    if n_gpus() < n:
        pytest.skip(f"{n_gpus()} GPU(s) on this box")
    r = subprocess.run([BIN, XML, str(n), "6"], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr[-2000:])
    if r.returncode == 77:
        pytest.skip(r.stdout.strip())
    assert r.returncode == 0, r.stderr[-3000:]
    assert f"ok on {n} GPU(s)" in r.stdout


def test_two_ranks_over_nccl(tmp_path):
    if n_gpus() < 2:
        pytest.skip(f"{n_gpus()} GPU(s) on this box")
    out = str(tmp_path / "verdict.json")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + os.getpid() % 300), os.path.join(ROOT, "tests", "mp_reduce_worker.py"), out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stderr[-4000:]
    v = json.load(open(out))
    print(v)
    assert v["world"] == 2 and v["nccl_ranks"] == 2
    assert all(v["per_rank_bit_exact"]), "a rank's frames differ from the same frames rendered on GPU 0"
    assert v["reduced_equals_sum_bits"] and v["host_buffers_equal_device"]      # two ranks: one summation order
    assert v["max_rel_vs_one_gpu"] <= 1e-5 and v["accumulated_max_rel_vs_one_gpu"] <= 1e-5
    assert v["counts"] == [6.0, 6.0, 6]
    assert v["non_root_host_buffers_untouched"]
