"""Builds libpupil_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(CSRC, "libpupil_b200.so")
SOURCES = ["pb_capi.cu", "pb_commit.cu", "pb_entry.cu", "pb_build.cu", "pb_trace.cu", "pb_wavefront.cu", "pb_comm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "--extended-lambda", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-Xptxas", "-v"]


def _newer(target: str, deps: list[str]) -> bool:
    return os.path.exists(target) and all(os.path.getmtime(target) >= os.path.getmtime(d) for d in deps)


def build_native(force: bool = False, verbose: bool = False) -> str:
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "pupil_b200.h"))
    if not force and _newer(OUT, deps):
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    env = dict(os.environ)
    env["PATH"] = "/usr/bin:" + env.get("PATH", "")   # host compiler: the system g++
    objs = []

    def compile_one(src):
        obj = os.path.join(CSRC, src[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("PB_NVCC_EXTRA", "").split(), "-ccbin", "/usr/bin/g++", "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True, env=env)
        log = r.stdout + r.stderr
        with open(os.path.join(CSRC, src[:-3] + ".ptxas.log"), "w") as f:
            f.write(log)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{log}")
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=6) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-ccbin", "/usr/bin/g++", "-o", OUT, *objs, "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return OUT


HOST_DIR = os.path.join(HERE, "host")
HOST_OUT = os.path.join(HOST_DIR, "libpupil_host.so")
HOST_SOURCES = ["scene_xml.cpp", "image_io.cpp", "b200_pt_pass.cpp", "host_capi.cpp"]


def build_host(force: bool = False) -> str:
    """g++ build of the C++ host layer (scene XML loader, headless System/Pass, B200PTPass) ->
    host/libpupil_host.so, linked against csrc/libpupil_b200.so ($ORIGIN rpath) and cudart."""
    build_native()
    deps = [os.path.join(HOST_DIR, f) for f in os.listdir(HOST_DIR) if f.endswith((".cpp", ".h"))]
    deps += [os.path.join(os.path.dirname(HERE), "include", h) for h in ("pupil_b200.h", "pupil_host.h")] + [OUT]
    if not force and _newer(HOST_OUT, deps):
        return HOST_OUT
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-Wno-unused-function", "-shared",
           "-I", os.path.join(cuda, "include"), *[os.path.join(HOST_DIR, f) for f in HOST_SOURCES], "-o", HOST_OUT,
           "-L", CSRC, "-lpupil_b200", "-L", os.path.join(cuda, "lib64"), "-lcudart", "-lz", "-Wl,-rpath,$ORIGIN/../csrc",
           "-Wl,-rpath," + os.path.join(cuda, "lib64")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("host layer build failed:\n" + r.stdout + r.stderr)
    return HOST_OUT


TESTS_DIR = os.path.join(os.path.dirname(HERE), "tests")


def build_devsim(force: bool = False) -> str:
    """TEST HARNESS (not product): tests/devsim/devsim.cu -> tests/devsim/libdevsim.so, the product's shading headers
    on the GPU outside the integrator, same nvcc flags as the product (-fmad=false).  Built here because the .so
    travels to the GPU box with the snapshot."""
    d = os.path.join(TESTS_DIR, "devsim")
    out = os.path.join(d, "libdevsim.so")
    deps = [os.path.join(d, f) for f in os.listdir(d) if f.endswith((".cu", ".h"))]
    deps += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    if not force and _newer(out, deps):
        return out
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    env = dict(os.environ)
    env["PATH"] = "/usr/bin:" + env.get("PATH", "")
    flags = [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    cmd = [nvcc, *flags, "-ccbin", "/usr/bin/g++", "-shared", os.path.join(d, "devsim.cu"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError("devsim build failed:\n" + r.stdout + r.stderr)
    return out


def build_cpp_tests(force: bool = False) -> str:
    """TEST PROGRAM (not product): tests/cpp/test_multi_gpu.cpp -> tests/cpp/test_multi_gpu, a plain C++ host of the two
    C ABIs (no Python, no torch).  Built here so that the binary travels to the GPU box."""
    build_host()
    d = os.path.join(TESTS_DIR, "cpp")
    out = os.path.join(d, "test_multi_gpu")
    src = os.path.join(d, "test_multi_gpu.cpp")
    deps = [src, HOST_OUT, OUT, os.path.join(os.path.dirname(HERE), "include", "pupil_host.h"), os.path.join(os.path.dirname(HERE), "include", "pupil_b200.h")]
    if not force and _newer(out, deps):
        return out
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(cuda, "include"), src, "-o", out,
           "-L", HOST_DIR, "-lpupil_host", "-L", CSRC, "-lpupil_b200", "-L", os.path.join(cuda, "lib64"), "-lcudart",
           "-Wl,-rpath,$ORIGIN/../../pupiloptixlab_b200/host", "-Wl,-rpath,$ORIGIN/../../pupiloptixlab_b200/csrc",
           "-Wl,-rpath," + os.path.join(cuda, "lib64")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("tests/cpp build failed:\n" + r.stdout + r.stderr)
    return out


if __name__ == "__main__":
    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
    print(build_devsim(force="--force" in sys.argv))
    print(build_cpp_tests(force="--force" in sys.argv))
