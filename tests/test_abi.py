"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares; ctypes mirrors match the C structs."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

from heatray_b200 import capi

ROOT = Path(__file__).resolve().parent.parent


def _declared(header):
    text = (ROOT / "include" / header).read_text()
    return re.findall(r"HRC_API\s+[\w\s\*]+?\b(hr[ch]\w+)\s*\(", text)


def test_backend_header_symbols_exported(lib):
    names = _declared("heatray_cuda.h")
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in heatray_cuda.h but not exported"
    assert sorted(n[3:] for n in names) == sorted(capi.BACKEND_SYMBOLS)


def test_host_header_symbols_exported(lib):
    names = _declared("heatray_host.h")
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in heatray_host.h but not exported"


def test_oracle_exports_same_abi(oracle_mod):
    o = oracle_mod.load()
    for n in capi.BACKEND_SYMBOLS:
        if n in ("GetDevicePixels", "TraceRaysNoAlloc") or n.startswith("Exchange"):          # device-memory entry points: nothing for a CPU restatement to mirror
            continue
        assert hasattr(o, "orc" + n), n


def test_struct_layouts_match_c(tmp_path):
    names = {"HrcContextDesc": capi.ContextDesc, "HrcTextureDesc": capi.TextureDesc, "HrcPbrParams": capi.PbrParams, "HrcGlassParams": capi.GlassParams,
             "HrcMeshDesc": capi.MeshDesc, "HrcEnvLight": capi.EnvLight, "HrcDirLight": capi.DirLight, "HrcPointLight": capi.PointLight,
             "HrcSpotLight": capi.SpotLight, "HrcGlobals": capi.Globals, "HrcFrameParams": capi.FrameParams, "HrcPassStats": capi.PassStats,
             "HrcSceneInfo": capi.SceneInfo, "HrcRay": None, "HrcHit": None, "HrhPbrMaterial": capi.HrhPbrMaterial, "HrhGlassMaterial": capi.HrhGlassMaterial}
    src = tmp_path / "s.c"
    src.write_text('#include <stdio.h>\n#include "heatray_host.h"\nint main(){' + "".join(f'printf("{n} %zu\\n", sizeof({n}));' for n in names) + "return 0;}")
    exe = tmp_path / "s"
    subprocess.run(["gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for n, t in names.items():
        size = C.sizeof(t) if t is not None else {"HrcRay": capi.RAY_DTYPE.itemsize, "HrcHit": capi.HIT_DTYPE.itemsize}[n]
        assert int(out[n]) == size, n
    assert int(out["HrcGlobals"]) == 84          # GlobalData, PassGenerator.h:269-294
    assert int(out["HrcRay"]) == 32 and int(out["HrcHit"]) == 16


def test_product_fails_loudly_without_gpu(lib):
    """No CPU fallback: without a device the backend refuses to create a context."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.HrcError) as e:
        capi.Backend(lib, "hrc")
    assert e.value.code == 6 and "no CPU path" in str(e.value)


def test_product_does_not_link_the_oracle():
    for lib_path in (capi.LIB_PATH, capi.PKG / "libheatray_host.so", capi.PKG / "libheatray_openrl.so"):
        out = subprocess.run(["nm", "-D", str(lib_path)], capture_output=True, text=True).stdout
        assert " orc" not in out and "liboracle" not in subprocess.run(["ldd", str(lib_path)], capture_output=True, text=True).stdout
    for f in (ROOT / "heatray_b200").rglob("*.py"):
        text = f.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle", text, re.M) and "liboracle" not in text, f
    for f in (ROOT / "heatray_b200" / "csrc").rglob("*"):
        if f.is_dir():
            continue
        assert "oracle/" not in f.read_text().replace("the CPU oracle", ""), f


def test_host_library_has_no_device_code_and_openrl_shim_exports_the_entry_points():
    """libheatray_host.so (what the CPU reference arm loads) carries no CUDA: no cudart symbols, no fatbin section; it reaches the backend only
    through weak references. libheatray_openrl.so exports every entry point include/heatray_openrl.h declares."""
    host = capi.PKG / "libheatray_host.so"
    syms = subprocess.run(["nm", "-D", str(host)], capture_output=True, text=True).stdout
    assert "cuda" not in syms.lower() and ".nv_fatbin" not in subprocess.run(["readelf", "-S", str(host)], capture_output=True, text=True).stdout
    assert all(line.split()[0] == "w" for line in syms.splitlines() if " hrc" in line and line.split()[-1].startswith("hrc"))
    assert "libheatray_host.so" in subprocess.run(["ldd", str(capi.LIB_PATH)], capture_output=True, text=True).stdout
    header = (ROOT / "include" / "heatray_openrl.h").read_text()
    wanted = set(re.findall(r"\b((?:rl[A-Z]|OpenRL|hrl)\w+)\s*\(", header)) - {"rlGetError"} | {"rlGetError"}
    exported = {line.split()[-1] for line in subprocess.run(["nm", "-D", "--defined-only", str(capi.PKG / "libheatray_openrl.so")], capture_output=True, text=True).stdout.splitlines() if " T " in line}
    assert len(wanted) >= 62 and wanted <= exported, sorted(wanted - exported)
