"""The drop-in boundary without a GPU: libpupil_b200.so loads, exports every PB_API symbol that
include/pupil_b200.h declares (and nothing of the oracle), its POD structs have the sizes the ctypes
mirror assumes, and it FAILS LOUDLY without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from pupiloptixlab_b200 import build as pb_build
from pupiloptixlab_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pupil_b200.h")


@pytest.fixture(scope="module")
def lib_path():
    return pb_build.build_native()


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"PB_API\s+[\w\s\*]+?\b(pb_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ("pb_create", "pb_destroy", "pb_scene_begin", "pb_add_material", "pb_add_trimesh", "pb_add_sphere", "pb_set_env",
              "pb_set_camera", "pb_scene_commit", "pb_update_instance", "pb_trace_closest", "pb_trace_any", "pb_render",
              "pb_render_ex", "pb_resolve", "pb_get_stats", "pb_comm_unique_id", "pb_comm_init_rank", "pb_create_multi", "pb_render_multi"):
        assert s in syms
    # every entry point cites the reference interface it replaces
    src = open(HEADER).read()
    assert src.count("framework/") + src.count("path_tracer/") >= 20


def test_library_exports_every_declared_symbol(lib_path):
    lib = C.CDLL(lib_path)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} is declared in include/pupil_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(pb_\w+)\b", out))
    assert set(declared_symbols()) <= exported
    assert "orc_" not in out, "the product must not contain oracle symbols"
    deps = subprocess.run(["ldd", lib_path], capture_output=True, text=True).stdout
    assert "liboracle" not in deps and "torch" not in deps
    assert "nccl" not in deps, "NCCL is bound at run time (dlopen), not linked: a second libnccl next to torch's must not be loaded"


def test_host_library_exports_every_declared_symbol():
    host_header = os.path.join(ROOT, "include", "pupil_host.h")
    declared = sorted(set(re.findall(r"PH_API\s+[\w\s\*]+?\b(ph_\w+)\s*\(", open(host_header).read())))
    assert {"ph_scene_load_xml", "ph_scene_upload", "ph_run_path_tracer", "ph_session_create", "ph_session_create_multi",
            "ph_session_move_object", "ph_session_read", "ph_image_load"} <= set(declared)
    so = pb_build.build_host()
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(ph_\w+)\b", out))
    assert set(declared) <= exported, sorted(set(declared) - exported)
    assert "orc_" not in out


def test_struct_sizes_match_the_header(lib_path, tmp_path):
    """sizeof() of the C structs (compiled from the header with gcc) == the ctypes mirror."""
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include "pupil_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(pb_texture), '
                    'sizeof(pb_material), sizeof(pb_env), sizeof(pb_ray), sizeof(pb_hit), sizeof(pb_render_params), sizeof(pb_stats), sizeof(pb_render_targets));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])   # the header is plain C
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    mirror = [C.sizeof(capi.Texture), C.sizeof(capi.Material), C.sizeof(capi.Env), capi.RAY_DTYPE.itemsize, capi.HIT_DTYPE.itemsize,
              C.sizeof(capi.RenderParams), C.sizeof(capi.Stats), C.sizeof(capi.RenderTargets)]
    assert sizes == mirror


def test_no_cpu_fallback(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    api = capi.load()
    assert "sm_100a" in api.version()
    with pytest.raises(capi.ApiError):
        capi.Context(api, 0)          # PB_ERR_NO_DEVICE: the product never computes on the CPU
    h = C.c_void_p()
    assert api._create(C.byref(h), 0) == 3
