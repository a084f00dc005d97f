#!/usr/bin/env python
"""Writes tests/golden/images_ref.npz: small PNG / JPEG / Radiance-HDR files (made here with Pillow and a few lines of
hand-rolled PNG chunk writing for the cases Pillow cannot produce) together with what the reference's decoder — stb_image.h as
vendored by the reference, compiled unmodified into oracle/_ref/libref_stb.so — returns for each of them. Run in the build
container after oracle/ref_probe/build_ref.sh; tests/test_images.py compares host_image.cpp's own decoders against these."""
import ctypes as C
import io
import struct
import sys
import zlib
from pathlib import Path

import numpy as np
from PIL import Image

ROOT = Path(__file__).resolve().parents[2]
REF = ROOT / "oracle" / "_ref" / "libref_stb.so"
OUT = ROOT / "tests" / "golden" / "images_ref.npz"


def stb_decode(lib, data: bytes, flip: bool):
    w, h, c = C.c_int(), C.c_int(), C.c_int()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    n = lib.ref_stb_load(buf, len(data), int(flip), C.byref(w), C.byref(h), C.byref(c), None, 0)
    if n == 0:
        raise RuntimeError("stb_image rejected a fixture")
    out = np.zeros(n, np.float32 if lib.ref_stb_is_hdr(buf, len(data)) else np.uint8)
    lib.ref_stb_load(buf, len(data), int(flip), C.byref(w), C.byref(h), C.byref(c), out.ctypes.data_as(C.c_void_p), n)
    return out.reshape(h.value, w.value, c.value)


def png_chunk(kind: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data) & 0xFFFFFFFF)


def raw_png(samples: np.ndarray, depth: int, color: int, interlace: bool, extra=b"") -> bytes:
    """samples: (h, w, comps) integer array of file samples (already in 0..2^depth-1); filter type 0 on every row."""
    h, w, comps = samples.shape

    def rows(sub):
        out = bytearray()
        for r in sub:
            flat = r.reshape(-1)
            if depth == 16:
                b = flat.astype(">u2").tobytes()
            elif depth == 8:
                b = flat.astype(np.uint8).tobytes()
            else:
                bits = np.zeros(((len(flat) * depth + 7) // 8) * 8, np.uint8)
                for k in range(depth):
                    bits[k:len(flat) * depth:depth] = (flat >> (depth - 1 - k)) & 1
                b = np.packbits(bits).tobytes()
            out += b"\x00" + b
        return bytes(out)
    if not interlace:
        raw = rows(samples)
    else:
        raw = b""
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = samples[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]:
                raw += rows(sub)
    ihdr = struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 1 if interlace else 0)
    return b"\x89PNG\r\n\x1a\n" + png_chunk(b"IHDR", ihdr) + extra + png_chunk(b"IDAT", zlib.compress(raw, 6)) + png_chunk(b"IEND", b"")


def rgbe_file(rgb: np.ndarray, rle: bool) -> bytes:
    h, w, _ = rgb.shape
    m = rgb.max(-1)
    e = np.where(m > 1e-32, np.floor(np.log2(np.maximum(m, 1e-38))) + 1, 0).astype(np.int32)
    scale = np.where(m > 1e-32, 256.0 / np.exp2(e.astype(np.float64)), 0.0)
    px = np.concatenate([np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8), np.where(m > 1e-32, e + 128, 0).astype(np.uint8)[..., None]], -1)
    body = bytearray()
    for y in range(h):
        if not rle:
            body += px[y].tobytes()
            continue
        body += bytes([2, 2, w >> 8, w & 255])
        for k in range(4):
            chan = px[y, :, k]
            i = 0
            while i < w:                       # runs of >= 4 as runs, everything else as dumps of <= 128
                run = 1
                while i + run < w and run < 127 and chan[i + run] == chan[i]:
                    run += 1
                if run >= 4:
                    body += bytes([128 + run, int(chan[i])])
                    i += run
                else:
                    j = i
                    while j < w and j - i < 128 and not (j + 3 < w and chan[j] == chan[j + 1] == chan[j + 2] == chan[j + 3]):
                        j += 1
                    j = max(j, i + 1)
                    body += bytes([j - i]) + chan[i:j].tobytes()
                    i = j
    return b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w) + bytes(body)


def main():
    if not REF.exists():
        sys.exit("oracle/_ref/libref_stb.so is missing: run oracle/ref_probe/build_ref.sh where /root/reference is mounted")
    lib = C.CDLL(str(REF))
    lib.ref_stb_load.restype = C.c_long
    rng = np.random.default_rng(20261104)
    h, w = 37, 53
    yy, xx = np.mgrid[0:h, 0:w]
    smooth = np.stack([xx * 4, yy * 6, (xx + yy) * 2], -1).astype(np.uint8)
    noise = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
    mixed = ((smooth.astype(np.int32) * 3 + noise) // 4).astype(np.uint8)
    files = {}

    def pil(name, img, fmt, **kw):
        bio = io.BytesIO()
        img.save(bio, fmt, **kw)
        files[name] = bio.getvalue()
    for nm, arr in (("smooth", smooth), ("mixed", mixed)):
        im = Image.fromarray(arr)
        pil(f"jpeg_{nm}_q90_420", im, "JPEG", quality=90)
        pil(f"jpeg_{nm}_q50_444", im, "JPEG", quality=50, subsampling=0)
        pil(f"jpeg_{nm}_q75_422", im, "JPEG", quality=75, subsampling=1)
        pil(f"jpeg_{nm}_q85_prog", im, "JPEG", quality=85, progressive=True)
        pil(f"jpeg_{nm}_q60_prog_420", im, "JPEG", quality=60, progressive=True, subsampling=2)
        pil(f"jpeg_{nm}_q80_opt", im, "JPEG", quality=80, optimize=True)
        pil(f"jpeg_{nm}_grey", im.convert("L"), "JPEG", quality=80)
        pil(f"jpeg_{nm}_grey_prog", im.convert("L"), "JPEG", quality=70, progressive=True)
        pil(f"jpeg_{nm}_restart", im, "JPEG", quality=80, restart_marker_blocks=2)
        for mode in ("RGB", "RGBA", "L", "LA", "P", "1"):
            pil(f"png_{nm}_{mode}", im.convert(mode), "PNG")
        pil(f"png_{nm}_P_trns", im.convert("P"), "PNG", transparency=3)
        pil(f"png_{nm}_L_trns", im.convert("L"), "PNG", transparency=int(arr[5, 5, 0]))
        pil(f"png_{nm}_RGB_trns", im, "PNG", transparency=tuple(int(v) for v in arr[7, 9]))
        pil(f"png_{nm}_P_4bit", im.convert("P", palette=Image.ADAPTIVE, colors=16), "PNG", bits=4)
        pil(f"png_{nm}_P_2bit", im.convert("P", palette=Image.ADAPTIVE, colors=4), "PNG", bits=2)
    pil("jpeg_odd_1x1", Image.fromarray(noise[:1, :1]), "JPEG", quality=90)
    pil("jpeg_odd_17x9_420", Image.fromarray(mixed[:9, :17]), "JPEG", quality=90, subsampling=2)
    pil("png_1x1", Image.fromarray(noise[:1, :1]), "PNG")
    big = (rng.integers(0, 65536, (h, w, 3))).astype(np.uint16)
    files["png_rgb16"] = raw_png(big, 16, 2, False)
    files["png_rgb16_adam7"] = raw_png(big, 16, 2, True)
    files["png_rgba8_adam7"] = raw_png(np.concatenate([mixed, noise[..., :1]], -1).astype(np.uint16), 8, 6, True)
    files["png_grey4_adam7"] = raw_png((mixed[..., :1] >> 4).astype(np.uint16), 4, 0, True)
    files["png_grey2"] = raw_png((mixed[..., :1] >> 6).astype(np.uint16), 2, 0, False)
    files["png_grey16_trns"] = raw_png(big[..., :1], 16, 0, False, extra=png_chunk(b"tRNS", struct.pack(">H", int(big[3, 4, 0]))))
    files["png_ga16"] = raw_png(big[..., :2], 16, 4, False)
    files["png_adam7_small_3x2"] = raw_png(mixed[:2, :3].astype(np.uint16), 8, 2, True)
    hdr = (rng.uniform(0, 1, (19, 40, 3)) ** 4 * 50).astype(np.float64)
    hdr[3:6, 5:30] = hdr[3, 5]                      # constant stretch: real runs in the RLE variant
    hdr[10, 0:8] = 0.0
    files["hdr_rle"] = rgbe_file(hdr, True)
    files["hdr_flat"] = rgbe_file(hdr, False)
    files["hdr_narrow_flat"] = rgbe_file(hdr[:, :5], False)

    out = {}
    for name, data in files.items():
        out["file/" + name] = np.frombuffer(data, np.uint8)
        out["ref/" + name] = stb_decode(lib, data, False)
        if name in ("png_mixed_RGBA", "jpeg_mixed_q90_420", "hdr_rle"):      # stbi_set_flip_vertically_on_load(true), as the reference loads textures
            out["ref_flipped/" + name] = stb_decode(lib, data, True)
    np.savez_compressed(OUT, **out)
    print(f"{len(files)} fixtures, {OUT.stat().st_size} bytes -> {OUT}")


if __name__ == "__main__":
    main()
