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


def _sources():
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp"))


def _headers():
    return sorted(CSRC.glob("*.h")) + sorted((ROOT / "include").glob("*.h")) + [Path(__file__)]      # the flags live in this file


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
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
            else:
                cmd += ["-x", "cu"] if False else []
            res = subprocess.run(cmd, capture_output=True, text=True)
            if verbose or res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed on {src.name}")
            (OBJ / (src.name + ".log")).write_text(res.stdout + res.stderr)
    if force or _stale(LIB, objs):
        cmd = [NVCC, *ARCH, "-shared", "-cudart", "static", "-o", str(LIB), *map(str, objs), "-Xcompiler", "-pthread"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("link failed")
    return LIB


APP = PKG / "heatray_headless"


def build_apps(force: bool = False) -> Path:
    """heatray_headless: the viewer's render loop without the viewer (csrc/apps/heatray_headless.cpp), linked against the in-tree library."""
    src = CSRC / "apps" / "heatray_headless.cpp"
    if force or _stale(APP, [src, LIB] + _headers()):
        cmd = ["g++", "-std=c++17", "-O2", "-pthread", str(src), "-o", str(APP), f"-L{PKG}", "-lheatray_cuda", "-Wl,-rpath,$ORIGIN"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("heatray_headless build failed")
    return APP


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_cuda(force, verbose)
    build_apps(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
