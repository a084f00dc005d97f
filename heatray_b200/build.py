"""In-tree build of libheatray_cuda.so (sm_100a only).

    python -m heatray_b200.build            # build everything that is out of date
    python -m heatray_b200.build --force

nvcc cross-compiles without a GPU; the built .so is git-ignored but travels to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
OBJ = PKG / "_build"
LIB = PKG / "libheatray_cuda.so"

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden,-ffp-contract=off,-fwrapv,-pthread"]
# -fmad=false: the shading arithmetic is evaluated operation by operation as the shader text (and the oracle, -ffp-contract=off) states it; FMAs
# appear only where the source writes fmaf / __fmaf_rn (traversal). Contraction was THE source of GPU-vs-oracle pixel differences: at 1080p pixels
# within 1e-4 of the oracle go from 99.0 % to 99.96 %, for 1.8 - 2.8 % of whole-pass throughput (profiles/r02_notes.md).
CU_FLAGS = ["--expt-relaxed-constexpr", "-fmad=false", "-Xptxas", "-v"]


HOST_LIB = PKG / "libheatray_host.so"
OPENRL_LIB = PKG / "libheatray_openrl.so"
# The host layer (QMC tables, material / light packing, mesh providers, image and scene readers, session files: no device code) is its own
# library, so that a process which only needs those (bench.py's CPU reference arm, the oracle-side renderer in the tests) never maps the CUDA
# backend. Its one backend caller, hrhSceneUpload, reaches the hrc* entry points through weak references that bind when libheatray_cuda.so
# is in the process (it links libheatray_host.so, so loading the backend brings both).
HOST_SOURCES = ["host_qmc.cpp", "host_scene.cpp", "host_image.cpp", "host_session.cpp", "host_import.cpp"]
SHIM_SOURCES = ["openrl_shim.cpp"]


def _sources():
    skip = set(HOST_SOURCES + SHIM_SOURCES)
    return [p for p in sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp")) if p.name not in skip]


def _headers():
    return sorted(CSRC.glob("*.h")) + sorted((ROOT / "include").glob("*.h")) + [Path(__file__)]      # the flags live in this file


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd, what):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(what)
    return res


def build_host(force: bool = False) -> Path:
    import re
    OBJ.mkdir(exist_ok=True)
    hdrs = _headers()
    weak = OBJ / "weak_backend.h"
    names = sorted(set(re.findall(r"\b(hrc[A-Z]\w+)\s*\(", (ROOT / "include" / "heatray_cuda.h").read_text())))
    text = '#include "../../include/heatray_cuda.h"\n' + "".join(f"#pragma weak {n}\n" for n in names)
    if not weak.exists() or weak.read_text() != text:
        weak.write_text(text)
    objs = []
    for name in HOST_SOURCES:
        src, obj = CSRC / name, OBJ / (name + ".o")
        objs.append(obj)
        if force or _stale(obj, [src, weak] + hdrs):
            _run(["g++", "-std=c++17", "-O3", "-fPIC", "-fvisibility=hidden", "-ffp-contract=off", "-fwrapv", "-pthread", "-include", str(weak), "-c", str(src), "-o", str(obj)], f"g++ failed on {name}")
    if force or _stale(HOST_LIB, objs):
        _run(["g++", "-shared", "-o", str(HOST_LIB), *map(str, objs), "-pthread"], "host library link failed")
    return HOST_LIB


def build_openrl(force: bool = False) -> Path:
    """libheatray_openrl.so: the OpenRL entry points over the backend C-ABI (include/heatray_openrl.h); finds libheatray_cuda.so next to itself at run time."""
    src = CSRC / "openrl_shim.cpp"
    if force or _stale(OPENRL_LIB, [src] + _headers()):
        _run(["g++", "-std=c++17", "-O2", "-fPIC", "-fvisibility=hidden", "-shared", str(src), "-o", str(OPENRL_LIB), "-ldl"], "libheatray_openrl.so build failed")
    return OPENRL_LIB


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    build_host(force)
    OBJ.mkdir(exist_ok=True)
    objs = []
    hdrs = _headers()
    for src in _sources():
        obj = OBJ / (src.name + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = [NVCC, *ARCH, *COMMON, "-c", str(src), "-o", str(obj)]
            if src.suffix == ".cu":
                cmd += CU_FLAGS
            res = subprocess.run(cmd, capture_output=True, text=True)
            if verbose or res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed on {src.name}")
            (OBJ / (src.name + ".log")).write_text(res.stdout + res.stderr)
    if force or _stale(LIB, objs + [HOST_LIB]):
        cmd = [NVCC, *ARCH, "-shared", "-cudart", "static", "-o", str(LIB), *map(str, objs), f"-L{PKG}", "-lheatray_host", "-Xlinker", "-rpath=$ORIGIN", "-Xcompiler", "-pthread"]
        _run(cmd, "link failed")
    return LIB


APP = PKG / "heatray_headless"


def build_apps(force: bool = False) -> Path:
    """heatray_headless: the viewer's render loop without the viewer (csrc/apps/heatray_headless.cpp), linked against the in-tree library."""
    src = CSRC / "apps" / "heatray_headless.cpp"
    if force or _stale(APP, [src, LIB] + _headers()):
        cmd = ["g++", "-std=c++17", "-O2", "-pthread", str(src), "-o", str(APP), f"-L{PKG}", "-lheatray_cuda", "-lheatray_host", "-Wl,-rpath,$ORIGIN"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("heatray_headless build failed")
    return APP


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_cuda(force, verbose)
    build_openrl(force)
    build_apps(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
