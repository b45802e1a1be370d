"""The PRODUCT's BSDF code (pupiloptixlab_b200/csrc/pb_bsdf.cuh + pb_shade.cuh + the host material prep of
pb_host_prep.h) against the per-BSDF known-answer tables produced by the REFERENCE's own headers
(tests/golden/reference_headers.npz `bsdf_*`, 15 parameterisations: the seven types, checkerboard parameters,
nonlinear plastics, a dielectric with int_ior < ext_ior, ...).  tests/golden pins the oracle's port; this file pins
pb_bsdf.cuh itself, twice:

* on the CPU (`-m "not gpu"`): the headers are PB_HD, tests/hostsim compiles them with g++ -ffp-contract=off and the
  same glibc sinf / cosf the golden generator used -> the 16-float records must be BIT-EXACT;
* on the GPU (`-m gpu`): the same body in a test-only kernel (tests/devsim/devsim.cu, the product's nvcc flags).
  libdevice sin / cos / sqrt differ from glibc by ulps, so a record matches when the sampled lobe and the RNG state
  after sampling are equal and every float is within 2e-4 relative (1e-6 absolute); an ulp can flip the
  reflect / refract or specular / diffuse choice, so >= 99.5 % of the records must match.
"""
import ctypes as C
import os

import numpy as np
import pytest

from pupiloptixlab_b200 import capi
from tests.golden import make_golden as G

GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_headers.npz")
N_BSDF = 1500   # make_golden.generate's n_bsdf


def _table(fn, mat, wo, seeds):
    out = np.zeros((len(wo), 16), np.float32)
    rc = fn(C.byref(mat), wo.ctypes.data_as(C.POINTER(C.c_float)), seeds.ctypes.data_as(C.POINTER(C.c_uint32)), len(wo),
            out.ctypes.data_as(C.POINTER(C.c_float)))
    assert rc == 0
    return out


def _cases():
    gold = np.load(GOLD)
    for k, (name, mat) in enumerate(G.golden_materials()):
        wo, seeds = G.bsdf_inputs(N_BSDF, 100 + k)
        yield name, mat, np.ascontiguousarray(wo), np.ascontiguousarray(seeds), gold[f"bsdf_{name}"]


def test_product_bsdf_headers_on_cpu_are_bit_exact(hostsim_lib):
    fn = hostsim_lib.hs_bsdf_table
    fn.restype = C.c_int
    for name, mat, wo, seeds, want in _cases():
        got = _table(fn, mat, wo, seeds)
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), f"{name}: {np.argwhere(~same)[:5].tolist()} differ, e.g. {got[~same][:3]} vs {want[~same][:3]}"


@pytest.mark.gpu
def test_product_bsdf_headers_on_gpu_match_reference_tables():
    so = os.path.join(os.path.dirname(__file__), "devsim", "libdevsim.so")
    assert os.path.exists(so), "tests/devsim/libdevsim.so is missing: run __graft_entry__.build()"
    fn = C.CDLL(so).ds_bsdf_table
    fn.restype = C.c_int
    worst = {}
    for name, mat, wo, seeds, want in _cases():
        got = _table(fn, mat, wo, seeds)
        lobe_ok = (got[:, 7].view(np.uint32) == want[:, 7].view(np.uint32)) & (got[:, 8].view(np.uint32) == want[:, 8].view(np.uint32))
        cols = [0, 1, 2, 3, 4, 5, 6, 9, 10, 11, 12]
        a, b = got[:, cols].astype(np.float64), want[:, cols].astype(np.float64)
        with np.errstate(invalid="ignore"):
            close = (np.abs(a - b) <= 1e-6 + 2e-4 * np.abs(b)) | (np.isnan(a) & np.isnan(b)) | ((a == b))
        ok = lobe_ok & close.all(axis=1)
        worst[name] = float(ok.mean())
        assert ok.mean() >= 0.995, f"{name}: only {ok.mean():.4f} of the records match"
    print("fraction of matching records per material:", worst)
