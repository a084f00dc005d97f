"""Host QMC generators (heatray_b200/csrc/host_qmc.cpp) against golden tables produced by the reference's own Random.h
(tests/golden/qmc_ref.json, written by oracle/ref_probe/make_golden.py) — bit-exact — and SURVEY.md Appendix C.1."""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = json.loads((Path(__file__).parent / "golden" / "qmc_ref.json").read_text())
u32, i32, f32 = C.c_uint32, C.c_int32, C.c_float


def _run(fn, n, *args):
    a = np.zeros((n, 2), np.float32)
    fn(a.ctypes.data_as(C.c_void_p), *args)
    return a


def _call(lib, kind, s, n):
    if kind == "sobol": return _run(lib.hrhSobol, n, u32(n), u32(s))
    if kind == "halton": return _run(lib.hrhHalton, n, u32(n), u32(s))
    if kind == "hammersley": return _run(lib.hrhHammersley, n, u32(n), u32(s))
    if kind == "radial_sobol": return _run(lib.hrhRadialSobol, n, u32(n), u32(s))
    if kind == "uniform": return _run(lib.hrhUniformRandom, n, u32(n), u32(s), f32(0.0), f32(1.0))
    if kind == "blue_noise": return _run(lib.hrhBlueNoise, n, u32(n), i32(s))
    if kind.startswith("polygonal"): return _run(lib.hrhRandomPolygonal, n, u32(int(kind[9:])), u32(n), u32(s))
    raise KeyError(kind)


@pytest.mark.parametrize("key", sorted(GOLD))
def test_generator_bit_exact_vs_reference(lib, key):
    kind, s, n = key.split("/")
    got = _call(lib, kind, int(s), int(n)).view(np.uint32).reshape(-1)
    assert np.array_equal(got, np.asarray(GOLD[key], np.uint32))


def test_survey_appendix_c1(lib):
    a = _call(lib, "sobol", 0, 8)
    assert np.allclose(a[:4], [[0.312489927, 0.743847013], [0.949272215, 0.30650726], [0.578496635, 0.94785428], [0.19015497, 0.192578599]], rtol=0, atol=5e-9)
    assert np.allclose(_call(lib, "sobol", 3, 8)[:2], [[0.858047843, 0.0507308356], [0.407982469, 0.831305742]], atol=5e-9)
    assert np.allclose(_call(lib, "halton", 1, 4), [[0.776115119, 0.167333245], [0.080465354, 0.0477572791], [0.704762995, 0.250755191], [0.271888226, 0.582632899]], atol=5e-9)
    assert np.allclose(_call(lib, "polygonal6", 0, 4), [[0.236698627, 0.422910243], [0.501299739, 0.546863675], [0.767811477, 0.333964944], [0.38221848, 0.670080781]], atol=5e-9)


def test_sobol_is_prefix_stable(lib):
    """What makes extending SequenceOffsets by H entries safe (quirk B-Q1)."""
    assert np.array_equal(_call(lib, "sobol", 0, 64)[:8], _call(lib, "sobol", 0, 8))


def test_live_reference_if_present(lib, oracle_mod):
    ref = oracle_mod.ref_lib("qmc")
    if ref is None:
        pytest.skip("oracle/_ref not built")
    for s in (2, 9):
        a = _run(lib.hrhSobol, 1000, u32(1000), u32(s)); b = _run(ref.ref_sobol, 1000, u32(1000), u32(s))
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_generate_random_sequences_layout(lib):
    """generateRandomSequences: sequence-major vec2[16*N], sequence i seeded by i; aperture table alongside (PassGenerator.cpp:603-684)."""
    n = 32
    xy, ap = np.zeros((16, n, 2), np.float32), np.zeros((16, n, 2), np.float32)
    lib.hrhGenerateRandomSequences(4, 0, 16, n, xy.ctypes.data_as(C.c_void_p), ap.ctypes.data_as(C.c_void_p))
    for s in (0, 5, 15):
        assert np.array_equal(xy[s], _call(lib, "sobol", s, n))
        assert np.array_equal(ap[s], _call(lib, "radial_sobol", s, n))
    lib.hrhGenerateRandomSequences(1, 2, 16, n, xy.ctypes.data_as(C.c_void_p), ap.ctypes.data_as(C.c_void_p))
    assert np.array_equal(xy[3], _call(lib, "halton", 3, n)) and np.array_equal(ap[3], _call(lib, "polygonal6", 3, n))
    assert (ap >= 0).all() and (ap <= 1).all()


def test_pixel_tables(lib):
    w, h = 48, 32
    n = lib.hrhSequenceOffsetCount(w, h)
    assert n >= w * h
    sq = lib.hrhSequenceOffsetCount(40, 40)
    assert sq == 40 * 40 + 20           # square frames index past w*h by H/2 (quirk B-Q1)
    off = np.zeros(n, np.float32)
    lib.hrhGenerateSequenceOffsets(w, h, off.ctypes.data_as(C.c_void_p), n)
    assert np.array_equal(off[:64], _call(lib, "sobol", 0, 64)[:, 0])
    sid = np.zeros(w * h, np.uint8)
    lib.hrhGenerateSequenceIds(w, h, 16, sid.ctypes.data_as(C.c_void_p))
    assert sid.max() <= 16 and len(np.unique(sid)) >= 12
    # perspective.rlsl:62 + utility.rlsl:15-18 in numpy fp32 (sin differs by ulps between libms, so allow a few pixels)
    x, y = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    d = ((x + np.float32(0.5)) / np.float32(w)) * np.float32(12.9898) + ((y + np.float32(0.5)) / np.float32(h)) * np.float32(78.233)
    v = np.sin(d.astype(np.float32)).astype(np.float32) * np.float32(43758.5453123)
    ref = np.floor((v - np.floor(v)) * np.float32(16)).astype(np.uint8).reshape(-1)
    assert (ref == sid).mean() > 0.98
