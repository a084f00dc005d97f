"""hrcTraceRaysNoAlloc (include/heatray_cuda.h): hrcTraceRays for callers that cannot call cudaMalloc — the device staging of rays and hits
is a static device array inside the library. No counterpart in the reference.

Checked through the C-ABI against the existing entry point, hrcTraceRays, on the same rays: hit / miss, primitive, back-face bit and
(t, u, v) must be bit-identical (the fixed-ray-set bar of tests/test_gpu_parity.py: both run the same kernel). Two calls in a row with
different ray sets and sizes: the array keeps the first call's contents, the second must not see them. This file sorts last on
purpose."""
import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

pytestmark = pytest.mark.gpu


def _rays(rng, n, spec):
    r = np.zeros(n, capi.RAY_DTYPE)
    r["origin"] = rng.uniform(-6.0, 6.0, (n, 3)).astype(np.float32) + np.array([0.0, 2.0, 0.0], np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    r["direction"] = d / np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    r["tmin"] = 0.0
    r["tmax"] = np.where(rng.random(n) < 0.2, rng.uniform(0.5, 5.0, n), np.inf).astype(np.float32)
    return r


def _identical(a, b):
    assert np.array_equal(a["primId"], b["primId"])
    for f in ("t", "u", "v"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


def test_no_alloc_trace_matches_trace_rays_on_two_calls_in_a_row(gpu_backend):
    b = gpu_backend
    r = renderer.Renderer(b)
    spec = scenes.sphere_grid(width=64, height=36, passes=4, grid=4, u_slices=24, v_slices=12, env_size=(64, 32))
    r.load_scene(spec)
    rng = np.random.default_rng(11)
    first, second = _rays(rng, 200_000, spec), _rays(rng, 70_001, spec)

    want_first = b.trace_rays(first)
    got_first = b.trace_rays_no_alloc(first)
    assert (want_first["primId"] >= 0).sum() > 1000 and (want_first["primId"] < 0).sum() > 1000
    _identical(got_first, want_first)

    got_second = b.trace_rays_no_alloc(second)                     # smaller set right after: the array still holds the first call's hits
    _identical(got_second, b.trace_rays(second))

    _identical(b.trace_rays_no_alloc(first, any_hit=True), b.trace_rays(first, any_hit=True))
    _identical(b.trace_rays_no_alloc(first[:1]), want_first[:1])
    assert len(b.trace_rays_no_alloc(first[:0])) == 0


def test_no_alloc_trace_rejects_more_than_the_maximum(gpu_backend):
    b = gpu_backend
    r = renderer.Renderer(b)
    r.load_scene(scenes.cornell_box(width=32, height=32, passes=4))
    rng = np.random.default_rng(5)
    full = _rays(rng, capi.NO_ALLOC_MAX_RAYS, None)
    _identical(b.trace_rays_no_alloc(full), b.trace_rays(full))     # exactly the maximum is fine
    with pytest.raises(capi.HrcError, match="at most HRC_NO_ALLOC_MAX_RAYS"):
        b.trace_rays_no_alloc(_rays(rng, capi.NO_ALLOC_MAX_RAYS + 1, None))
    _identical(b.trace_rays_no_alloc(full[:1000]), b.trace_rays(full[:1000]))   # the context is still usable after the rejection
