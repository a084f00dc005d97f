"""Texture decoding and screenshot encoding (heatray_b200/csrc/host_image.cpp; reference Source/Utility/TextureLoader.cpp:17-159,
HeatrayRenderer.cpp:1619-1662). The decoders are pinned bit for bit against stb_image.h as vendored by the reference (compiled
unmodified into oracle/_ref/libref_stb.so by oracle/ref_probe/build_ref.sh; fixtures + expected texels in tests/golden/images_ref.npz,
written by oracle/ref_probe/make_image_golden.py)."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi

GOLD = Path(__file__).resolve().parent / "golden" / "images_ref.npz"


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _names(g):
    return sorted(k[5:] for k in g.files if k.startswith("file/"))


def test_every_fixture_decodes_like_the_reference_decoder(lib, gold):
    names = _names(gold)
    assert len(names) >= 50 and any(n.startswith("jpeg_") and "prog" in n for n in names) and any("adam7" in n for n in names)
    for n in names:
        got = capi.load_image(gold["file/" + n].tobytes())
        want = gold["ref/" + n]
        assert got.dtype == want.dtype and got.shape == want.shape, (n, got.shape, want.shape)
        assert np.array_equal(got, want), n          # bit-exact: integer reconstruction for PNG/JPEG, mantissa * 2^(e-136) for RGBE
    for k in (k for k in gold.files if k.startswith("ref_flipped/")):
        n = k[len("ref_flipped/"):]
        assert np.array_equal(capi.load_image(gold["file/" + n].tobytes(), flip=True), gold[k]), n


def test_live_stb_reference_if_present(lib, oracle_mod):
    """Where oracle/_ref is built and the reference is mounted: the reference's own image files through both decoders."""
    ref = Path(oracle_mod.HERE) / "_ref" / "libref_stb.so"
    files = [Path("/root/reference/Resources/Environments/studio.hdr"), Path("/root/reference/Resources/logo.png"), Path("/root/reference/ExampleImages/example2.png")]
    if not ref.exists() or not all(f.exists() for f in files):
        pytest.skip("reference not mounted here")
    stb = C.CDLL(str(ref))
    stb.ref_stb_load.restype = C.c_long
    for f in files:
        data = f.read_bytes()
        w, h, c = C.c_int(), C.c_int(), C.c_int()
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        n = stb.ref_stb_load(buf, len(data), 1, C.byref(w), C.byref(h), C.byref(c), None, 0)
        want = np.zeros(n, np.float32 if f.suffix == ".hdr" else np.uint8)
        stb.ref_stb_load(buf, len(data), 1, C.byref(w), C.byref(h), C.byref(c), want.ctypes.data_as(C.c_void_p), n)
        got = capi.load_image(f, flip=True)
        assert got.shape == (h.value, w.value, c.value) and np.array_equal(got.reshape(-1), want), f


def test_corrupt_and_unknown_files_fail_loudly(lib, gold, tmp_path):
    with pytest.raises(capi.HrcError):
        capi.load_image(b"not an image at all")
    png = gold["file/png_mixed_RGB"].tobytes()
    with pytest.raises(capi.HrcError):
        capi.load_image(png[:40])
    with pytest.raises(capi.HrcError):
        capi.load_image(tmp_path / "missing.png")
    with pytest.raises(capi.HrcError):
        capi.load_texture(tmp_path / "missing.png")


TIFF_GOLD = Path(__file__).resolve().parent / "golden" / "tiff_ref.npz"


def test_tiff_fixtures_and_the_reference_lut_file(lib, tmp_path):
    """Baseline TIFF (the FreeImage branch of util::loadTexture, TextureLoader.cpp:44-87). Expected texels come from Pillow, an independent
    decoder (oracle/ref_probe/make_tiff_golden.py); one fixture is the reference's own Resources/multiscatter_lut.tiff."""
    g = np.load(TIFF_GOLD)
    names = sorted(k[5:] for k in g.files if k.startswith("file/"))
    assert "reference_multiscatter_lut" in names and len(names) >= 11
    for n in names:
        got, want = capi.load_image(g["file/" + n].tobytes()), g["ref/" + n]
        assert got.dtype == want.dtype and got.shape == want.shape, n
        assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), n
    # the texture loader on a .tiff: FreeImage bitmaps are bottom-up, one float channel is RL_LUMINANCE (:69-72), default sampler clamps
    path = tmp_path / "lut.tiff"
    path.write_bytes(g["file/reference_multiscatter_lut"].tobytes())
    px, fmt, ws, wt = capi.load_texture(path, generate_mips=False, convert_to_linear=False)
    assert (fmt, ws, wt) == (capi.TEX_R32F, capi.WRAP_CLAMP, capi.WRAP_CLAMP)
    assert np.array_equal(px[..., 0], g["ref/reference_multiscatter_lut"][::-1, :, 0])
    # unsupported flavours are refused: 16-bit samples, 8-bit RGB (the reference would upload it as RGBA)
    from PIL import Image
    import io
    for arr in (np.zeros((4, 4), np.uint16), np.zeros((4, 4, 3), np.uint8)):
        b = io.BytesIO()
        Image.fromarray(arr).save(b, format="TIFF")
        with pytest.raises(capi.HrcError):
            capi.load_image(b.getvalue())


def _mutate(rng, data, structural=b""):
    b = bytearray(data)
    for _ in range(int(rng.integers(1, 5))):
        if len(b) <= 4:
            break
        pos = int(rng.integers(0, len(b)))
        op = int(rng.integers(0, 5 if structural else 4))
        if op == 0:
            b[pos] = int(rng.integers(0, 256))
        elif op == 1:
            del b[pos:pos + int(rng.integers(1, 30))]
        elif op == 2:
            b[pos:pos] = bytes(rng.integers(0, 256, int(rng.integers(1, 8)), dtype=np.uint8))
        elif op == 3:
            del b[max(4, pos):]
        else:
            b[pos] = structural[int(rng.integers(0, len(structural)))]
    return bytes(b)


def test_malformed_images_are_rejected_not_crashed_on(lib, gold):
    """Seeded mutations of every fixture (bit flips, cuts, insertions, truncation) plus three hand-made cases that used to crash or
    over-allocate: the decoders must answer with an error code or an image, never take the process down."""
    rng = np.random.default_rng(2024)
    files = [k for k in gold.files if k.startswith("file/")]
    tg = np.load(TIFF_GOLD)
    blobs = [gold[k].tobytes() for k in files] + [tg[k].tobytes() for k in tg.files if k.startswith("file/")]
    decoded = rejected = 0
    for t in range(1500):
        data = _mutate(rng, blobs[t % len(blobs)])
        try:
            im = capi.load_image(data)
            assert im.size > 0
            decoded += 1
        except capi.HrcError:
            rejected += 1
    assert decoded > 50 and rejected > 50
    # (a) a JPEG whose scan names Huffman tables that no DHT segment defined
    jpg = bytearray(gold[[k for k in files if "jpeg" in k][0]].tobytes())
    out, i = bytearray(jpg[:2]), 2
    while i + 4 <= len(jpg) and jpg[i] == 0xFF and jpg[i + 1] != 0xDA:
        seg = 2 + (jpg[i + 2] << 8 | jpg[i + 3])
        if jpg[i + 1] != 0xC4:
            out += jpg[i:i + seg]
        i += seg
    out += jpg[i:]
    with pytest.raises(capi.HrcError):
        capi.load_image(bytes(out))
    # (b) a PNG header that claims 2^24 x 2^24 pixels
    png = bytearray(gold[[k for k in files if "png" in k][0]].tobytes())
    png[16:24] = (1 << 24).to_bytes(4, "big") * 2
    with pytest.raises(capi.HrcError):
        capi.load_image(bytes(png))
    # (c) a Radiance header with an absurd resolution line
    with pytest.raises(capi.HrcError):
        capi.load_image(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 2000000000 +X 2000000000\n" + bytes(64))


def test_load_texture_follows_the_reference_loader(lib, gold, tmp_path):
    """util::loadTextureInternal: vertical flip, sRGB -> linear requantisation with truncation and alpha left alone
    (TextureLoader.cpp:109-133), format from the channel count (:138-151), sampler state (:37-42)."""
    p = tmp_path / "t.png"
    p.write_bytes(gold["file/png_mixed_RGBA"].tobytes())
    src = gold["ref_flipped/png_mixed_RGBA"]
    px, fmt, ws, wt = capi.load_texture(p, generate_mips=True, convert_to_linear=True)
    assert fmt == capi.TEX_RGBA8 and (ws, wt) == (capi.WRAP_REPEAT, capi.WRAP_REPEAT)
    v = src[..., :3].astype(np.float32) / np.float32(255.0)
    lin = np.where(v <= np.float32(0.04045), v / np.float32(12.92), np.power((v + np.float32(0.055)) / np.float32(1.055), np.float32(2.4), dtype=np.float32))
    want = (lin * np.float32(255.0)).astype(np.uint8)
    assert np.abs(px[..., :3].astype(int) - want.astype(int)).max() <= 1 and (px[..., :3] == want).mean() > 0.999   # numpy pow vs powf: a truncation boundary may differ
    assert np.array_equal(px[..., 3], src[..., 3])
    px2, fmt2, ws2, wt2 = capi.load_texture(p, generate_mips=False, convert_to_linear=False)
    assert np.array_equal(px2, src) and (ws2, wt2) == (capi.WRAP_CLAMP, capi.WRAP_CLAMP)
    # luminance replicates into RGB with alpha 1 (RL_LUMINANCE), RGB gets alpha 255
    g = tmp_path / "g.png"
    g.write_bytes(gold["file/png_mixed_L"].tobytes())
    pg, fg, _, _ = capi.load_texture(g, True, False)
    want_g = np.flipud(gold["ref/png_mixed_L"])
    assert fg == capi.TEX_RGBA8 and all(np.array_equal(pg[..., k], want_g[..., 0]) for k in range(3)) and (pg[..., 3] == 255).all()
    # two-channel images are skipped by the reference (TextureLoader.cpp:146-151): an error here, never a guess
    la = tmp_path / "la.png"
    la.write_bytes(gold["file/png_mixed_LA"].tobytes())
    with pytest.raises(capi.HrcError):
        capi.load_texture(la)
    # the "../" fallback (:24-35)
    sub = tmp_path / "sub"
    sub.mkdir()
    import os
    cwd = os.getcwd()
    try:
        os.chdir(sub)
        px3, *_ = capi.load_texture("t.png", False, False)
        assert np.array_equal(px3, src)
    finally:
        os.chdir(cwd)
    # Radiance .hdr comes back as float RGB
    hp = tmp_path / "e.hdr"
    hp.write_bytes(gold["file/hdr_rle"].tobytes())
    ph, fh, _, _ = capi.load_texture(hp, True, True)
    assert fh == capi.TEX_RGB32F and np.array_equal(ph, gold["ref_flipped/hdr_rle"])


def test_writers_round_trip(lib, tmp_path):
    rng = np.random.default_rng(5)
    for shape in ((9, 13, 3), (300, 257, 3), (4, 4, 4), (7, 5, 1), (3, 2, 2)):       # 300x257x3 spans several 64 KiB stored blocks
        a = rng.integers(0, 256, shape).astype(np.uint8)
        capi.write_png(tmp_path / "o.png", a, rows_bottom_up=True)
        back = capi.load_image(tmp_path / "o.png")
        assert np.array_equal(back, np.flipud(a))
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(tmp_path / "o.png")).reshape(shape), np.flipud(a))    # an independent decoder accepts the file
    f = (rng.uniform(0, 1, (11, 17, 3)) ** 3 * 40).astype(np.float32)
    f[2, 3] = 0.0
    capi.write_hdr(tmp_path / "o.hdr", f, rows_bottom_up=True)
    back = capi.load_image(tmp_path / "o.hdr")
    assert back.shape == f.shape
    top = np.flipud(f)
    assert (np.abs(back - top) <= top.max(-1, keepdims=True) / 128 + 1e-6).all()        # 8-bit mantissas under one shared exponent: one step of the brightest channel
    capi.write_hdr(tmp_path / "o.pfm", f, rows_bottom_up=True, pfm=True)
    raw = (tmp_path / "o.pfm").read_bytes()
    head = b"PF\n17 11\n-1.0\n"
    assert raw.startswith(head) and np.array_equal(np.frombuffer(raw[len(head):], "<f4").reshape(11, 17, 3), f)
