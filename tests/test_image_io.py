"""Host image I/O (pupiloptixlab_b200/host/image_io.cpp) = util::BitmapTexture::Load / Save of the reference
(framework/util/texture.cpp:13-180, which wraps stb_image / stb_image_write / tinyexr).  Files are produced here
from the format specifications with zlib / struct (no image library), independently of the C++ writers."""
import os
import struct
import zlib

import numpy as np
import pytest

from pupiloptixlab_b200 import hostlib


def _chunk(tag, body):
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def write_png(path, img, ctype, depth=8, palette=None, trns=None, filters=(0, 1, 2, 3, 4)):
    """img: (h, w, c) integer samples; rows use the filter types in turn (all five reconstruction paths)."""
    h, w, c = img.shape
    bpp = max(1, c * depth // 8)
    rows = []
    prev = np.zeros(0, np.uint8)
    for y in range(h):
        if depth == 8:
            line = img[y].astype(np.uint8).reshape(-1)
        elif depth == 16:
            line = img[y].astype(">u2").reshape(-1).view(np.uint8)
        else:
            bits = np.unpackbits(img[y].astype(np.uint8).reshape(-1, 1), axis=1)[:, 8 - depth:].reshape(-1)
            line = np.packbits(bits)
        line = line.astype(np.int32)
        up = prev if len(prev) else np.zeros_like(line)
        ft = filters[y % len(filters)]
        left = np.concatenate([np.zeros(bpp, np.int32), line[:-bpp]])
        ul = np.concatenate([np.zeros(bpp, np.int32), up[:-bpp]])
        if ft == 0: f = line
        elif ft == 1: f = line - left
        elif ft == 2: f = line - up
        elif ft == 3: f = line - ((left + up) >> 1)
        else:
            p = left + up - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - up), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, up, ul))
            f = line - pred
        rows.append(bytes([ft]) + (f & 255).astype(np.uint8).tobytes())
        prev = line
    data = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    if palette is not None:
        data += _chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        data += _chunk(b"tRNS", np.asarray(trns, np.uint8).tobytes())
    z = zlib.compress(b"".join(rows))
    data += _chunk(b"IDAT", z[: len(z) // 2]) + _chunk(b"IDAT", z[len(z) // 2:]) + _chunk(b"IEND", b"")
    open(path, "wb").write(data)


def decode_ldr(u8, alpha=None):
    """texture.cpp:110-115"""
    rgb = np.power(u8.astype(np.float32) / np.float32(255), np.float32(2.2), dtype=np.float32)
    a = np.ones(u8.shape[:2] + (1,), np.float32) if alpha is None else (alpha.astype(np.float32) / np.float32(255))[..., None]
    return np.concatenate([rgb, a], -1)


def test_png_colour_types(tmp_path):
    rng = np.random.RandomState(0)
    rgb = rng.randint(0, 256, (13, 17, 3))
    write_png(tmp_path / "rgb.png", rgb, 2)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "rgb.png")), decode_ldr(rgb), rtol=2e-6, atol=1e-7)
    rgba = rng.randint(0, 256, (9, 11, 4))
    write_png(tmp_path / "rgba.png", rgba, 6)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "rgba.png")), decode_ldr(rgba[..., :3], rgba[..., 3]), rtol=2e-6, atol=1e-7)
    rgb16 = rng.randint(0, 65536, (5, 7, 3))
    write_png(tmp_path / "rgb16.png", rgb16, 2, depth=16)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "rgb16.png")), decode_ldr(rgb16 >> 8), rtol=2e-6, atol=1e-7)
    pal = rng.randint(0, 256, (16, 3))
    idx = rng.randint(0, 16, (6, 10, 1))
    write_png(tmp_path / "pal4.png", idx, 3, depth=4, palette=pal, trns=list(range(0, 160, 10)))
    got = hostlib.load_image(str(tmp_path / "pal4.png"))
    np.testing.assert_allclose(got, decode_ldr(pal[idx[..., 0]], (np.arange(0, 160, 10))[idx[..., 0]]), rtol=2e-6, atol=1e-7)


def test_grey_png_reproduces_the_reference_channel_walk(tmp_path):
    """c == 1: StbImageLoad reads data[i], data[i+1], data[i+2] with i += c (texture.cpp:110-116), i.e. the two
    following pixels end up in G and B.  Reproduced (clamped at the end of the buffer)."""
    g = np.random.RandomState(1).randint(0, 256, (4, 6, 1))
    write_png(tmp_path / "g.png", g, 0)
    got = hostlib.load_image(str(tmp_path / "g.png"))
    flat = g.reshape(-1)
    idx = np.minimum(np.arange(flat.size)[:, None] + np.arange(3)[None], flat.size - 1)
    np.testing.assert_allclose(got[..., :3].reshape(-1, 3), decode_ldr(flat[idx][None])[0, :, :3], rtol=2e-6, atol=1e-7)


def rgbe_rle_file(img):
    """Radiance .hdr with new-style run-length scanlines (Greg Ward's format)."""
    h, w, _ = img.shape
    mx = img.max(-1)
    m, e = np.frexp(mx)
    scale = np.where(mx > 1e-32, m * 256.0 / np.maximum(mx, 1e-38), 0)
    rgbe = np.zeros((h, w, 4), np.uint8)
    rgbe[..., :3] = (img * scale[..., None]).astype(np.uint8)
    rgbe[..., 3] = np.where(mx > 1e-32, e + 128, 0)
    out = b"#?RADIANCE\n# made by the test\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {h} +X {w}\n".encode()
    for y in range(h):
        out += bytes([2, 2, w >> 8, w & 255])
        for k in range(4):
            line = rgbe[y, :, k]
            x = 0
            while x < w:
                run = 1
                while x + run < w and run < 127 and line[x + run] == line[x]:
                    run += 1
                if run >= 3:
                    out += bytes([128 + run, line[x]]); x += run
                else:
                    n = min(w - x, 5)
                    out += bytes([n]) + line[x:x + n].tobytes(); x += n
    return out, rgbe


def test_hdr_rle_and_roundtrip(tmp_path):
    rng = np.random.RandomState(2)
    img = (rng.gamma(0.5, 4.0, (7, 12, 3)) * (rng.rand(7, 12, 1) > 0.2)).astype(np.float32)
    img[2, 3:9] = img[2, 3]            # runs
    data, rgbe = rgbe_rle_file(img)
    open(tmp_path / "a.hdr", "wb").write(data)
    got = hostlib.load_image(str(tmp_path / "a.hdr"))
    want = rgbe[..., :3].astype(np.float32) * np.ldexp(np.float32(1), rgbe[..., 3].astype(np.int32) - 136)[..., None]
    want[rgbe[..., 3] == 0] = 0
    np.testing.assert_array_equal(got[..., :3], want.astype(np.float32))
    assert (got[..., 3] == 1).all()
    # our writer (SaveHdr: flipped vertically, texture.cpp:14) read back
    rgba = np.concatenate([img, np.ones((7, 12, 1), np.float32)], -1)
    hostlib.save_image(str(tmp_path / "b.hdr"), rgba)
    back = hostlib.load_image(str(tmp_path / "b.hdr"))
    assert (np.abs(back[::-1, :, :3] - img) <= img.max(-1, keepdims=True) / 128 + 1e-12).all()     # 8-bit mantissa shared per pixel


def exr_file(img, compression, half=False):
    """Scanline OpenEXR written from the specification: channels A B G R (alphabetical), zip / zips / none."""
    h, w, c = img.shape
    names = ["A", "B", "G", "R"][4 - c:] if c < 4 else ["A", "B", "G", "R"]
    src = {"R": 0, "G": 1, "B": 2, "A": 3}
    head = struct.pack("<II", 20000630, 2)
    ch = b"".join(n.encode() + b"\0" + struct.pack("<IIII", 1 if half else 2, 0, 1, 1) for n in names) + b"\0"
    def attr(n, t, v): return n.encode() + b"\0" + t.encode() + b"\0" + struct.pack("<I", len(v)) + v
    head += attr("channels", "chlist", ch) + attr("compression", "compression", bytes([compression]))
    head += attr("dataWindow", "box2i", struct.pack("<iiii", 0, 0, w - 1, h - 1)) + attr("displayWindow", "box2i", struct.pack("<iiii", 0, 0, w - 1, h - 1))
    head += attr("lineOrder", "lineOrder", b"\0") + attr("pixelAspectRatio", "float", struct.pack("<f", 1)) + b"\0"
    lines = 16 if compression == 3 else 1
    chunks = []
    for y0 in range(0, h, lines):
        raw = b""
        for y in range(y0, min(h, y0 + lines)):
            for n in names:
                raw += img[y, :, src[n]].astype("<f2" if half else "<f4").tobytes()
        if compression:
            a = np.frombuffer(raw, np.uint8)
            t = np.concatenate([a[0::2], a[1::2]]).astype(np.int32)
            d = t.copy(); d[1:] = (t[1:] - t[:-1] + 128 + 256) & 255
            z = zlib.compress(d.astype(np.uint8).tobytes())
            body = z if len(z) < len(raw) else raw
        else:
            body = raw
        chunks.append(struct.pack("<iI", y0, len(body)) + body)
    off = len(head) + 8 * len(chunks)
    table = b""
    for cnk in chunks:
        table += struct.pack("<Q", off); off += len(cnk)
    return head + table + b"".join(chunks)


@pytest.mark.parametrize("compression,half,chans", [(0, False, 3), (2, False, 4), (3, False, 3), (3, True, 4)])
def test_exr_variants(tmp_path, compression, half, chans):
    rng = np.random.RandomState(3)
    img = rng.gamma(0.7, 2.0, (37, 23, chans)).astype(np.float32)
    img[5:20, 3:15] = 0.25          # compressible area (exercises the deflated branch)
    if half:
        img = img.astype(np.float16).astype(np.float32)
    p = tmp_path / "a.exr"
    open(p, "wb").write(exr_file(img, compression, half))
    got = hostlib.load_image(str(p))
    np.testing.assert_array_equal(got[..., :3], img[..., :3])
    np.testing.assert_array_equal(got[..., 3], img[..., 3] if chans == 4 else np.ones((37, 23), np.float32))


def test_exr_and_pfm_writers_roundtrip(tmp_path):
    rng = np.random.RandomState(4)
    rgba = rng.gamma(0.7, 2.0, (11, 19, 4)).astype(np.float32)
    hostlib.save_image(str(tmp_path / "o.exr"), rgba)
    back = hostlib.load_image(str(tmp_path / "o.exr"))
    np.testing.assert_array_equal(back[::-1, :, :3], rgba[..., :3])      # SaveExr writes the rows bottom-up (texture.cpp:36-44)
    raw = open(tmp_path / "o.exr", "rb").read()
    assert raw.index(b"B\0") < raw.index(b"G\0") < raw.index(b"R\0")      # B, G, R channel order
    hostlib.save_image(str(tmp_path / "o.pfm"), rgba)
    np.testing.assert_array_equal(hostlib.load_image(str(tmp_path / "o.pfm"))[..., :3], rgba[..., :3])
    u8 = rng.randint(0, 256, (6, 5, 4)).astype(np.uint8)
    hostlib.save_image(str(tmp_path / "o.png"), u8)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "o.png")), decode_ldr(u8[..., :3], u8[..., 3]), rtol=2e-6, atol=1e-7)


def test_bmp_ppm_and_errors(tmp_path):
    rng = np.random.RandomState(5)
    rgb = rng.randint(0, 256, (5, 7, 3)).astype(np.uint8)
    open(tmp_path / "a.ppm", "wb").write(b"P6\n# c\n7 5\n255\n" + rgb.tobytes())
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "a.ppm")), decode_ldr(rgb), rtol=2e-6, atol=1e-7)
    stride = (7 * 3 + 3) & ~3
    rows = b"".join(rgb[y, :, ::-1].tobytes() + b"\0" * (stride - 21) for y in range(4, -1, -1))
    bmp = b"BM" + struct.pack("<IHHI", 54 + len(rows), 0, 0, 54) + struct.pack("<IiiHHIIiiII", 40, 7, 5, 1, 24, 0, len(rows), 0, 0, 0, 0) + rows
    open(tmp_path / "a.bmp", "wb").write(bmp)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "a.bmp")), decode_ldr(rgb), rtol=2e-6, atol=1e-7)
    open(tmp_path / "a.jpg", "wb").write(b"\xff\xd8\xff\xe0" + b"\0" * 32)
    with pytest.raises(hostlib.HostError, match="JPEG"):
        hostlib.load_image(str(tmp_path / "a.jpg"))
    with pytest.raises(hostlib.HostError):
        hostlib.load_image(str(tmp_path / "missing.png"))


def test_png_adam7(tmp_path):
    """Interlaced PNG: the seven passes written by hand."""
    rng = np.random.RandomState(6)
    h, w = 11, 13
    rgb = rng.randint(0, 256, (h, w, 3))
    passes = [(0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)]
    raw = b""
    for x0, y0, dx, dy in passes:
        sub = rgb[y0::dy, x0::dx]
        if sub.size == 0:
            continue
        for row in sub:
            raw += b"\0" + row.astype(np.uint8).tobytes()
    data = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 1)) + _chunk(b"IDAT", zlib.compress(raw)) + _chunk(b"IEND", b"")
    open(tmp_path / "i.png", "wb").write(data)
    np.testing.assert_allclose(hostlib.load_image(str(tmp_path / "i.png")), decode_ldr(rgb), rtol=2e-6, atol=1e-7)


XML_TEXTURED = """<scene version="3.0.0">
  <integrator type="path"><integer name="max_depth" value="5"/></integrator>
  <sensor type="perspective">
    <float name="fov" value="40"/>
    <transform name="to_world"><lookat origin="0, 1, 4" target="0, 0.5, 0" up="0, 1, 0"/></transform>
    <film type="hdrfilm"><integer name="width" value="64"/><integer name="height" value="48"/></film>
  </sensor>
  <bsdf type="diffuse" id="tex">
    <texture name="reflectance" type="bitmap">
      <string name="filename" value="tex/albedo.png"/>
      <string name="filter_type" value="bilinear"/>
      <string name="wrap_mode" value="mirror"/>
      <transform name="to_uv"><scale x="3" y="2"/></transform>
    </texture>
  </bsdf>
  <shape type="rectangle">
    <transform name="to_world"><rotate x="1" angle="-90"/><scale value="2"/></transform>
    <ref id="tex"/>
  </shape>
  <shape type="sphere"><float name="radius" value="0.5"/><point name="center" x="0" y="0.5" z="0"/>
    <bsdf type="roughconductor"><float name="alpha" value="0.2"/></bsdf>
  </shape>
  <emitter type="envmap">
    <string name="filename" value="sky.exr"/>
    <float name="scale" value="2.5"/>
    <transform name="to_world"><rotate y="1" angle="30"/></transform>
  </emitter>
</scene>
"""


def test_xml_bitmap_texture_and_envmap(tmp_path, oracle_api):
    """scene.cpp:141-165 (bitmap texture: filename relative to the scene, filter_type, wrap_mode, to_uv) and
    :208-220 (envmap emitter: scale, filename, transform) end to end: XML + image files -> sink -> render, against
    the same scene described through the ABI with the arrays the files were made from."""
    from oracle import pyoracle
    from pupiloptixlab_b200 import capi, scenes
    os.makedirs(tmp_path / "tex")
    rng = np.random.RandomState(7)
    albedo = rng.randint(30, 256, (8, 12, 3))
    write_png(tmp_path / "tex" / "albedo.png", albedo, 2)
    sky = scenes.procedural_envmap(32, 16)
    open(tmp_path / "sky.exr", "wb").write(exr_file(sky[..., :3], 3))
    open(tmp_path / "scene.xml", "w").write(XML_TEXTURED)
    sc = hostlib.Scene(path=str(tmp_path / "scene.xml"))
    i = sc.info()
    assert (i.n_shapes, i.has_env, i.film_w, i.film_h) == (2, capi.ENV_MAP, 64, 48)   # has_env carries the env type
    t = sc.shape(0)["material"].reflectance
    assert (t.type, t.bitmap_w, t.bitmap_h, t.address_mode, t.filter_mode) == (capi.TEX_BITMAP, 12, 8, capi.ADDRESS_MIRROR, capi.FILTER_LINEAR)
    assert (t.uv_r0[0], t.uv_r1[1]) == (3, 2)
    np.testing.assert_allclose(np.ctypeslib.as_array(t.bitmap, shape=(8, 12, 4)), decode_ldr(albedo), rtol=2e-6, atol=1e-7)
    a = pyoracle.OracleContext(oracle_api)
    sc.upload(a)
    ra = a.render(64, 48, 5, 0, 4)[0]
    assert np.isfinite(ra).all() and ra[..., :3].mean() > 0.05 and a.stats()["n_emitters"] == 1
    # the same scene through the ABI
    b = pyoracle.OracleContext(oracle_api)
    b.scene_begin()
    M = capi.make_material
    tex = capi.bitmap_texture(decode_ldr(albedo), capi.ADDRESS_MIRROR, capi.FILTER_LINEAR, 3, 2)
    m0 = b.add_material(M(capi.MAT_DIFFUSE, reflectance=tex))
    s0, s1 = sc.shape(0), sc.shape(1)
    scenes.add_rectangle(b, np.vstack([s0["to_world"], [0, 0, 0, 1]]).astype(np.float32), m0, 0, None)
    b.add_sphere(np.asarray(s1["to_world"], np.float32), 0, b.add_material(s1["material"]), None)
    c30, s30 = np.cos(np.radians(30.0)), np.sin(np.radians(30.0))
    b.set_env(capi.env_map(sky, scale=2.5, to_world=[[c30, 0, s30, 0], [0, 1, 0, 0], [-s30, 0, c30, 0]]))
    s2c, c2w = sc.camera()
    b.set_camera(s2c, c2w)
    b.commit()
    rb = b.render(64, 48, 5, 0, 4)[0]
    err = np.abs(ra - rb)[..., :3] / (np.abs(rb[..., :3]) + 1e-3)
    assert (err.max(-1) <= 1e-3).mean() >= 0.99


def test_crafted_headers_are_refused_not_read_past_the_end(tmp_path):
    """the decoders run on untrusted texture files.  Explicit regressions next to the ASan/UBSan fuzzer
    (tests/test_sanitize_host.py): an 8-bit BMP whose palette would lie beyond the end of the file, a BMP with a pixel
    offset past the end, a PFM that ends inside its header, PNM / PFM / EXR headers with absurd dimensions."""
    def refused(name, data):
        p = tmp_path / name
        open(p, "wb").write(data)
        with pytest.raises(hostlib.HostError):
            hostlib.load_image(str(p))
    info = struct.pack("<IiiHHIIiiII", 40, 4, 2, 1, 8, 0, 8, 0, 0, 0, 0)
    refused("pal.bmp", b"BM" + struct.pack("<IHHI", 62, 0, 0, 54) + info + bytes(8))                 # 8 bpp, no palette in the file
    refused("hdr.bmp", b"BM" + struct.pack("<IHHI", 62, 0, 0, 54) + struct.pack("<IiiHHIIiiII", 0x7fffff00, 4, 2, 1, 8, 0, 8, 0, 0, 0, 0) + bytes(8))
    refused("off.bmp", b"BM" + struct.pack("<IHHI", 62, 0, 0, 0xfffffff0) + struct.pack("<IiiHHIIiiII", 40, 4, 2, 1, 24, 0, 8, 0, 0, 0, 0) + bytes(32))
    refused("neg.bmp", b"BM" + struct.pack("<IHHI", 62, 0, 0, 54) + struct.pack("<IiiHHIIiiII", 40, 4, -2147483648, 1, 24, 0, 8, 0, 0, 0, 0) + bytes(32))
    refused("short.pfm", b"PF\n1")
    refused("tiny.pfm", b"PF")
    refused("huge.pfm", b"PF\n99999999 99999999\n-1.0\n" + bytes(64))
    refused("huge.ppm", b"P6\n99999999999999999999 5\n255\n" + bytes(64))
    refused("wide.ppm", b"P6\n70000 70000\n255\n" + bytes(64))
    good = T_exr_small()
    bad = bytearray(good)
    k = bad.index(b"dataWindow") + len(b"dataWindow\0box2i\0") + 4
    bad[k + 8:k + 12] = struct.pack("<i", 0x7ffffff0)                                                # xMax
    refused("window.exr", bytes(bad))


def T_exr_small():
    return exr_file(np.random.RandomState(1).rand(3, 4, 3).astype(np.float32), 0)
