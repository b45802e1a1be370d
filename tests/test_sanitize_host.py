"""ASan + UBSan over the host layer's parsers of untrusted files (SURVEY §5 sanitizer targets; ADVICE r1 on
image_io.cpp): tests/sanitize/fuzz_host.cpp is compiled with -fsanitize=address,undefined -fno-sanitize-recover from the
product's own sources (image_io.cpp, scene_xml.cpp — no CUDA, no library of this repo linked) and run over seed files of
every supported format plus ~10^4 deterministic mutations each (all truncation lengths, byte flips, extreme header words,
huge ASCII numbers).  A decoder may refuse a file; a sanitizer report fails the test."""
import os
import shutil
import struct
import subprocess
import zlib

import numpy as np
import pytest

from tests import test_image_io as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "pupiloptixlab_b200", "host")


@pytest.fixture(scope="module")
def fuzz_bin(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("san") / "fuzz_host")
    cmd = ["/usr/bin/g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-fno-omit-frame-pointer",
           "-I", "/usr/local/cuda/include", os.path.join(ROOT, "tests", "sanitize", "fuzz_host.cpp"), os.path.join(HOST, "image_io.cpp"),
           os.path.join(HOST, "scene_xml.cpp"), "-lz", "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        if "asan" in r.stderr.lower() and "cannot find" in r.stderr.lower():
            pytest.skip("libasan is not installed")
        raise AssertionError(r.stderr[-3000:])
    return out


def write_seeds(d):
    rng = np.random.RandomState(3)
    rgb = rng.randint(0, 256, (6, 9, 3)).astype(np.uint8)
    T.write_png(os.path.join(d, "rgb.png"), rgb, 2)
    T.write_png(os.path.join(d, "rgba.png"), rng.randint(0, 256, (5, 4, 4)), 6)
    T.write_png(os.path.join(d, "pal.png"), rng.randint(0, 16, (7, 8, 1)), 3, depth=4, palette=rng.randint(0, 256, (16, 3)), trns=list(range(0, 160, 10)))
    T.write_png(os.path.join(d, "g16.png"), rng.randint(0, 65536, (4, 5, 1)), 0, depth=16)
    raw = b"".join(b"\0" + rgb[y].tobytes() for y in range(6))
    open(os.path.join(d, "plain.ppm"), "wb").write(b"P6\n# c\n9 6\n255\n" + rgb.tobytes())
    open(os.path.join(d, "grey.pgm"), "wb").write(b"P5 9 6 100\n" + (rgb[..., 0] % 101).tobytes())
    stride = (9 * 3 + 3) & ~3
    rows = b"".join(rgb[y, :, ::-1].tobytes() + b"\0" * (stride - 27) for y in range(5, -1, -1))
    open(os.path.join(d, "a24.bmp"), "wb").write(b"BM" + struct.pack("<IHHI", 54 + len(rows), 0, 0, 54) + struct.pack("<IiiHHIIiiII", 40, 9, 6, 1, 24, 0, len(rows), 0, 0, 0, 0) + rows)
    pal = rng.randint(0, 256, (256, 4)).astype(np.uint8).tobytes()
    idx = rng.randint(0, 256, (6, 12)).astype(np.uint8)
    open(os.path.join(d, "a8.bmp"), "wb").write(b"BM" + struct.pack("<IHHI", 54 + 1024 + 72, 0, 0, 54 + 1024) + struct.pack("<IiiHHIIiiII", 40, 12, 6, 1, 8, 0, 72, 0, 0, 256, 0) + pal + idx.tobytes())
    tga_hdr = lambda typ, bpp, desc=0x20: struct.pack("<BBBHHBHHHHBB", 0, 0, typ, 0, 0, 0, 0, 0, 9, 6, bpp, desc)
    open(os.path.join(d, "raw.tga"), "wb").write(tga_hdr(2, 24) + rgb[..., ::-1].tobytes())
    rle = b"".join(bytes([0x80 | 8]) + bytes(rgb[y, 0, ::-1]) + bytes([0]) + bytes(rgb[y, 8, ::-1]) for y in range(6))
    open(os.path.join(d, "rle.tga"), "wb").write(tga_hdr(10, 24, 0) + rle)
    img = rng.gamma(0.6, 2.0, (5, 12, 3)).astype(np.float32)
    open(os.path.join(d, "rle.hdr"), "wb").write(T.rgbe_rle_file(img)[0])
    open(os.path.join(d, "flat.hdr"), "wb").write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 2 +X 3\n" + bytes(range(24)))
    for comp in (0, 2, 3):
        open(os.path.join(d, f"c{comp}.exr"), "wb").write(T.exr_file(img, comp))
    open(os.path.join(d, "half.exr"), "wb").write(T.exr_file(img, 3, half=True))
    open(os.path.join(d, "a.pfm"), "wb").write(b"PF\n4 3\n-1.0\n" + rng.rand(3, 4, 3).astype("<f4").tobytes())
    open(os.path.join(d, "g.pfm"), "wb").write(b"Pf\n4 3\n1.0\n" + rng.rand(3, 4).astype(">f4").tobytes())
    for name in ("cornell_box.xml", "cornell_materials.xml", "ref_material_test.xml", "ref_mis.xml"):
        shutil.copy(os.path.join(ROOT, "tests", "data", name), os.path.join(d, name))
    os.makedirs(os.path.join(d, "tex"), exist_ok=True)
    T.write_png(os.path.join(d, "tex", "albedo.png"), rgb, 2)
    open(os.path.join(d, "sky.exr"), "wb").write(T.exr_file(img, 3))
    open(os.path.join(d, "textured.xml"), "w").write(T.XML_TEXTURED)


def test_image_decoders_and_scene_loader_under_asan_ubsan(fuzz_bin, tmp_path):
    write_seeds(str(tmp_path))
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0:allocator_may_return_null=1:max_allocation_size_mb=2048", UBSAN_OPTIONS="print_stacktrace=1")
    r = subprocess.run([fuzz_bin, str(tmp_path)], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-4000:])
    assert "no sanitizer report" in r.stdout
    print(r.stdout.strip())
