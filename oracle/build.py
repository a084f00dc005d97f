"""Builds the CPU oracle (test infrastructure): oracle/liboracle*.so and, where /root/reference is mounted, oracle/_ref/*.so."""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def build_oracle(force: bool = False) -> None:
    odir = ROOT / "oracle"
    args = ["make", "-C", str(odir)]
    if force:
        args.append("-B")
    res = subprocess.run(args, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("oracle build failed")
    ref = odir / "ref_probe" / "build_ref.sh"
    if Path("/root/reference/Source").is_dir() and ref.exists():
        have = all((odir / "_ref" / n).exists() for n in ("libref_qmc.so", "libref_lut.so", "libref_rlsl.so", "libref_display.so", "libref_stb.so", "libref_session.so", "libref_shaders.so", "libref_scene.so"))
        if force or not have:
            res = subprocess.run(["sh", str(ref)], capture_output=True, text=True)
            if res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
                raise RuntimeError("oracle/_ref build failed")


if __name__ == "__main__":
    build_oracle(force="--force" in sys.argv)
