"""Loads the CPU oracle (TEST INFRASTRUCTURE). Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
reference arm may import this module; the product package never does."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent


def _ensure_built():
    if not (HERE / "liboracle.so").exists() or not (HERE / "liboracle_fast.so").exists():
        subprocess.run(["make", "-C", str(HERE)], check=True, capture_output=True)


def load(fast: bool = False) -> C.CDLL:
    _ensure_built()
    lib = C.CDLL(str(HERE / ("liboracle_fast.so" if fast else "liboracle.so")))
    lib.orcGetCapturedRays.restype = C.c_uint64
    lib.orcGetCapturedRays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    lib.orcSetRayCapture.argtypes = [C.c_void_p, C.c_uint64]
    lib.orcSetThreads.argtypes = [C.c_void_p, C.c_int32]
    lib.orcSetBruteForce.argtypes = [C.c_void_p, C.c_int32]
    return lib


def backend(fast: bool = False, threads: int = 0):
    """A capi.Backend bound to the oracle's `orc*` exports."""
    from heatray_b200 import capi
    b = capi.Backend(load(fast), prefix="orc")
    if threads:
        b.lib.orcSetThreads(b.ctx, threads)
    return b


def capture_rays(b, max_rays: int):
    b.lib.orcSetRayCapture(b.ctx, max_rays)


def captured_rays(b):
    from heatray_b200 import capi
    n = b.lib.orcGetCapturedRays(b.ctx, None, None, 0)
    rays = np.zeros(n, capi.RAY_DTYPE)
    kinds = np.zeros(n, np.int32)
    b.lib.orcGetCapturedRays(b.ctx, rays.ctypes.data_as(C.c_void_p), kinds.ctypes.data_as(C.c_void_p), n)
    return rays, kinds


def ref_lib(name: str):
    """oracle/_ref/libref_<name>.so built from the reference's own sources (None when absent)."""
    p = HERE / "_ref" / f"libref_{name}.so"
    return C.CDLL(str(p)) if p.exists() else None
