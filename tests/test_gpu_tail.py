"""The tail of a pass (k_tail_decide / k_path, HRC_TUNE_TAIL_RAYS): once an iteration's extension queue is small the remaining paths are
followed to their end by ONE kernel instead of by the wavefront's per-iteration launches. It calls the same device functions in the
same per-pixel order, so the accumulation buffer, the ray statistics and the per-pixel ray counts must be BIT-IDENTICAL to the pure
wavefront's — for any threshold, including one so large that the whole frame is 'tail' from the primary rays on."""
import os

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

pytestmark = pytest.mark.gpu


def _render(lib, spec, tail_rays, passes=3, overlap=None, ray_counts=False, mutate=None):
    old = {k: os.environ.get(k) for k in ("HRC_TUNE_TAIL_RAYS", "HRC_TUNE_OVERLAP")}
    os.environ["HRC_TUNE_TAIL_RAYS"] = str(tail_rays)
    if overlap is not None:
        os.environ["HRC_TUNE_OVERLAP"] = str(overlap)
    try:
        b = capi.Backend(lib, "hrc", device=0, collect_stats=False)          # the tuning knobs are read when the context is created
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    try:
        r = renderer.Renderer(b)
        r.load_scene(spec)
        if mutate:
            mutate(b, r)
        if ray_counts:
            b.set_option(capi.OPT_PIXEL_RAY_COUNTS, 1)
        b.reset_stats()
        for i in range(passes):
            r.render_pass(want_stats=False, sample_index=i)
        st = b.get_stats()
        out = {"pixels": r.pixels().copy(), "rays": (int(st.raysPrimary), int(st.raysExtension), int(st.raysOcclusion))}
        if ray_counts:
            out["counts"] = b.ray_counts().copy()
        return out
    finally:
        b.close()


def _same(a, b):
    assert a["rays"] == b["rays"], (a["rays"], b["rays"])
    assert np.array_equal(a["pixels"], b["pixels"], equal_nan=True), int((a["pixels"] != b["pixels"]).any(-1).sum())
    if "counts" in a:
        assert np.array_equal(a["counts"], b["counts"])


SCENES = {
    "sphere_grid": lambda: scenes.sphere_grid(width=320, height=180, passes=8, grid=4, u_slices=24, v_slices=12, env_size=(128, 64), max_depth=8),
    "cornell": lambda: scenes.cornell_box(width=160, height=160, passes=8, max_depth=5),                       # punctual lights, closed box
    "textured_showcase": lambda: scenes.textured_showcase(width=192, height=128, passes=8, max_depth=6),      # glass, textures, alpha masks
    "material_grid": lambda: scenes.material_grid(width=240, height=135, passes=8, grid=5, u_slices=16, v_slices=8, max_depth=8),
}


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("tail_rays", [2000, 49152, 1 << 22])
def test_tail_is_bit_identical_to_the_wavefront(lib, name, tail_rays):
    spec = SCENES[name]()
    wavefront = _render(lib, spec, 0, ray_counts=True)
    assert wavefront["rays"][1] > 0
    _same(_render(lib, spec, tail_rays, ray_counts=True), wavefront)


def test_tail_without_stream_overlap(lib):
    spec = SCENES["textured_showcase"]()
    _same(_render(lib, spec, 49152, overlap=0), _render(lib, spec, 0, overlap=0))


def test_tail_with_bands_and_pipelined_passes(lib):
    """Enqueued passes rotate over the lanes, each with its own control words; a band renders a subset of the rows."""
    spec = SCENES["sphere_grid"]()

    def run(tail):
        old = os.environ.get("HRC_TUNE_TAIL_RAYS")
        os.environ["HRC_TUNE_TAIL_RAYS"] = str(tail)
        try:
            b = capi.Backend(lib, "hrc", device=0, collect_stats=False)
        finally:
            if old is None:
                os.environ.pop("HRC_TUNE_TAIL_RAYS", None)
            else:
                os.environ["HRC_TUNE_TAIL_RAYS"] = old
        try:
            r = renderer.Renderer(b)
            r.load_scene(spec)
            r.set_band(8, 1, 3)
            for i in range(8):
                r.enqueue_pass(sample_index=i)
            b.synchronize()
            st = b.get_stats()
            return {"pixels": r.pixels().copy(), "rays": (int(st.raysPrimary), int(st.raysExtension), int(st.raysOcclusion))}
        finally:
            b.close()

    a, w = run(49152), run(0)
    assert a["rays"] == w["rays"]
    # lanes fold their partial sums in a fixed order: same order with and without the tail
    assert np.array_equal(a["pixels"], w["pixels"])


def test_tail_shallow_and_deep_limits(lib):
    """maxRayDepth 0 (no extension rays at all) and 1 (the tail starts at the last iteration)."""
    for depth in (0, 1, 2):
        spec = scenes.sphere_grid(width=160, height=90, passes=4, grid=3, u_slices=16, v_slices=8, env_size=(64, 32), max_depth=depth)
        _same(_render(lib, spec, 1 << 22), _render(lib, spec, 0))
