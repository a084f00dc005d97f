#!/usr/bin/env python
"""Writes tests/golden/tiff_ref.npz: the reference's own Resources/multiscatter_lut.tiff (bytes) with its texels as decoded by Pillow —
an independent decoder — plus small TIFF fixtures covering what host_image.cpp's decoder accepts (float L / RGB, RGBA8; none / LZW /
PackBits; predictor 2; big-endian). Run in the build container, where /root/reference is mounted:
    python oracle/ref_probe/make_tiff_golden.py"""
import io
import struct
from pathlib import Path

import numpy as np
from PIL import Image

ROOT = Path(__file__).resolve().parents[2]
REFERENCE = Path("/root/reference/Resources/multiscatter_lut.tiff")
OUT = ROOT / "tests" / "golden" / "tiff_ref.npz"


def raw_tiff(pixels: np.ndarray, big_endian: bool, sample_format: int) -> bytes:
    """Uncompressed single-strip TIFF written by hand (Pillow cannot write float RGB or big-endian files)."""
    h, w, spp = pixels.shape
    bo = ">" if big_endian else "<"
    data = pixels.astype(pixels.dtype.newbyteorder(bo)).tobytes()
    bps = pixels.dtype.itemsize * 8
    entries = [(256, 3, 1, w), (257, 3, 1, h), (259, 3, 1, 1), (262, 3, 1, 2 if spp >= 3 else 1), (277, 3, 1, spp), (278, 3, 1, h),
               (279, 4, 1, len(data)), (284, 3, 1, 1)]
    extra = b""
    base = 8 + 2 + 12 * (len(entries) + 3) + 4

    def arr(tag, vals):
        nonlocal extra
        if len(vals) <= 2:
            packed = struct.pack(bo + "H" * len(vals), *vals).ljust(4, b"\0")
            return (tag, 3, len(vals), packed)
        off = base + len(extra)
        extra += struct.pack(bo + "H" * len(vals), *vals)
        return (tag, 3, len(vals), off)
    entries.append(arr(258, [bps] * spp))
    entries.append(arr(339, [sample_format] * spp))
    entries.append((273, 4, 1, None))
    body = b""
    strip_off = base + 0          # patched below once `extra` is final
    entries.sort(key=lambda e: e[0])
    strip_off = base + len(extra)
    for tag, typ, cnt, val in entries:
        if val is None:
            val = strip_off
        if isinstance(val, bytes):
            body += struct.pack(bo + "HHI", tag, typ, cnt) + val
        elif typ == 3 and cnt <= 2:
            body += struct.pack(bo + "HHI", tag, typ, cnt) + struct.pack(bo + "H", val).ljust(4, b"\0")
        else:
            body += struct.pack(bo + "HHII", tag, typ, cnt, val)
    head = (b"MM" + struct.pack(">H", 42) if big_endian else b"II" + struct.pack("<H", 42)) + struct.pack(bo + "I", 8)
    return head + struct.pack(bo + "H", len(entries)) + body + struct.pack(bo + "I", 0) + extra + data


def main():
    rng = np.random.default_rng(7)
    out = {}
    lut_bytes = REFERENCE.read_bytes()
    out["file/reference_multiscatter_lut"] = np.frombuffer(lut_bytes, np.uint8)
    out["ref/reference_multiscatter_lut"] = np.array(Image.open(io.BytesIO(lut_bytes)), np.float32)[..., None]
    lum = rng.normal(size=(19, 23)).astype(np.float32)
    rgba = rng.integers(0, 256, (21, 17, 4)).astype(np.uint8)
    smooth = np.clip(np.add.outer(np.arange(40), np.arange(36))[..., None] * np.array([1, 2, 3, 4]) + rng.integers(0, 3, (40, 36, 4)), 0, 255).astype(np.uint8)
    for comp in ("raw", "tiff_lzw", "packbits"):
        b = io.BytesIO(); Image.fromarray(lum).save(b, format="TIFF", compression=comp)
        out[f"file/float_l_{comp}"] = np.frombuffer(b.getvalue(), np.uint8); out[f"ref/float_l_{comp}"] = lum[..., None]
        b = io.BytesIO(); Image.fromarray(rgba, "RGBA").save(b, format="TIFF", compression=comp)
        out[f"file/rgba8_{comp}"] = np.frombuffer(b.getvalue(), np.uint8); out[f"ref/rgba8_{comp}"] = rgba[..., [2, 1, 0, 3]]     # FreeImage byte order
    b = io.BytesIO(); Image.fromarray(smooth, "RGBA").save(b, format="TIFF", compression="tiff_lzw", tiffinfo={317: 2})
    assert np.array_equal(np.array(Image.open(io.BytesIO(b.getvalue()))), smooth)
    out["file/rgba8_lzw_predictor2"] = np.frombuffer(b.getvalue(), np.uint8); out["ref/rgba8_lzw_predictor2"] = smooth[..., [2, 1, 0, 3]]
    rgbf = rng.uniform(0, 8, (9, 13, 3)).astype(np.float32)
    out["file/float_rgb_raw"] = np.frombuffer(raw_tiff(rgbf, False, 3), np.uint8); out["ref/float_rgb_raw"] = rgbf
    out["file/float_rgb_bigendian"] = np.frombuffer(raw_tiff(rgbf, True, 3), np.uint8); out["ref/float_rgb_bigendian"] = rgbf
    out["file/float_l_bigendian"] = np.frombuffer(raw_tiff(lum[..., None], True, 3), np.uint8); out["ref/float_l_bigendian"] = lum[..., None]
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, len(out) // 2, "fixtures")


if __name__ == "__main__":
    main()
