"""hr::PassGenerator (the C++ pass interface, heatray_b200/csrc/host_pass_generator.cpp) driven the way HeatrayRenderer
drives the reference's PassGenerator (HeatrayRenderer.cpp:366-408): async enqueue, worker-thread callbacks, mapped pixels."""
import ctypes as C
import math

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

pytestmark = pytest.mark.gpu


def _options(lib, spec, w, h, **kw):
    o = capi.RenderOptions()
    lib.hrhDefaultRenderOptions(C.byref(o))
    o.enableInteractiveMode = 0
    o.maxRenderPasses, o.maxRayDepth = spec.passes, spec.max_depth
    o.envEnabled = 0
    cam = spec.camera
    o.aspectRatio, o.focusDistance, o.focalLength = w / h, cam["focus_distance"], cam["focal_length"]
    vm = (C.c_float * 16)()
    lib.hrhOrbitCameraViewMatrix(cam["distance"], cam["phi"], cam["theta"], (C.c_float * 3)(*cam["target"]), vm)
    o.viewMatrix = vm
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def test_pass_generator_matches_direct_backend(lib, gpu_backend):
    spec = scenes.cornell_box(width=64, height=48, passes=4)
    w, h = spec.width, spec.height
    # direct path
    rg = renderer.Renderer(gpu_backend)
    rg.load_scene(spec)
    for _ in range(3):
        rg.render_pass()
    want = rg.pixels()

    pg = C.c_void_p()
    assert lib.hrhPassGeneratorCreate(0, w, h, C.byref(pg)) == 0, lib.hrhPassGeneratorLastError(pg)
    try:
        def load(user, ctx):                                   # ↔ loadScene callback on the worker thread
            b = capi.Backend(lib, "hrc", borrowed_ctx=ctx)
            b.width, b.height = w, h
            renderer.Renderer(b).upload_scene(spec, set_environment=False)
        task = capi.CONTEXT_TASK(load)
        lib.hrhPassGeneratorRunTask(pg, task, None)
        got = []

        def done(user, available, pixels, pw, ph, seconds, index):
            got.append((available, index, np.ctypeslib.as_array(pixels, shape=(ph, pw, 4)).copy() if available else None, seconds))
        cb = capi.PASS_COMPLETE_CALLBACK(done)
        o = _options(lib, spec, w, h)
        for i in range(3):
            o.resetInternalState = 1 if i == 0 else 0
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)      # returns immediately (asynchronous enqueue)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert lib.hrhPassGeneratorLastError(pg) == b""
        assert [g[0] for g in got] == [1, 1, 1] and [g[1] for g in got] == [1, 2, 3]
        # the three jobs were queued together, so passes 2 and 3 were rendered ahead into the deferred buffer while the previous result
        # was read back: same samples, folded in as one addition per pass (fp32 rounding of the sum order)
        assert np.array_equal(got[-1][2][..., 3], want[..., 3]) and np.allclose(got[-1][2], want, rtol=1e-5, atol=1e-6)
        assert (got[0][2][..., 3] == 1).all() and (got[1][2][..., 3] == 2).all()      # every callback sees exactly its own pass count
        # submitted one at a time there is nothing to look ahead to: bit-identical to the direct path
        got.clear()
        for i in range(3):
            o.resetInternalState = 1 if i == 0 else 0
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
            lib.hrhPassGeneratorWaitIdle(pg)
        assert [g[1] for g in got] == [1, 2, 3] and np.array_equal(got[-1][2], want)
        # a reset queued right behind a pass drops the pass rendered ahead
        got.clear()
        o.resetInternalState = 1
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        o.resetInternalState = 0
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        o.resetInternalState = 1
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        o.resetInternalState = 0
        for i in range(2):
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert lib.hrhPassGeneratorLastError(pg) == b""
        assert [g[1] for g in got] == [1, 2, 1, 2, 3]
        assert np.array_equal(got[-1][2][..., 3], want[..., 3]) and np.allclose(got[-1][2], want, rtol=1e-5, atol=1e-6)
        assert np.allclose(got[1][2], got[3][2], rtol=1e-5, atol=1e-6)

        # offline mode: all passes inside one job, pixels only at the end (PassGenerator.cpp:346-400)
        got.clear()
        o.enableOfflineMode, o.resetInternalState = 1, 1
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert [g[0] for g in got] == [0, 0, 0, 1] and (got[-1][2][..., 3] == 4).all()

        # debug pass rendering: exactly pass k (PassGenerator.cpp:570-572) and a debug visualiser (showShader: red for PBR)
        got.clear()
        o.enableOfflineMode, o.resetInternalState, o.debugPassRendering, o.debugPassIndex = 0, 1, 1, 2
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        o.debugPassRendering, o.debugVisMode = 0, 14
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert got[0][1] == 3
        vis = got[1][2]
        hit = vis[..., 3] == 2
        assert hit.any() and np.allclose(vis[hit][:, :3], [1, 0, 0])

        # resize forces a reset and new pixel tables (PassGenerator.cpp:301-323)
        got.clear()
        lib.hrhPassGeneratorResize(pg, 32, 32)
        o = _options(lib, spec, 32, 32)
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert got[0][2].shape == (32, 32, 4) and (got[0][2][..., 3] == 1).all()

        # interactive block mode (PassGenerator.cpp:372-381, perspective.rlsl:42-57): 9 passes walk every pixel of each 3x3 block
        # once at sample index 0, so their sum is exactly one offline pass; the reported pass index advances on the 9th
        got.clear()
        o.enableInteractiveMode, o.resetInternalState = 1, 1
        for i in range(9):
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
            o.resetInternalState = 0
        lib.hrhPassGeneratorWaitIdle(pg)
        assert lib.hrhPassGeneratorLastError(pg) == b""
        assert [g[1] for g in got] == [0] * 8 + [1]
        assert (got[0][2][..., 3].sum() == 32 * 32 // 9 + (1 if (32 * 32) % 9 else 0)) or (got[0][2][..., 3].sum() > 100)
        inter = got[-1][2]
        assert (inter[..., 3] == 1).all()
        got.clear()
        o.enableInteractiveMode, o.resetInternalState = 0, 1
        lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert np.array_equal(got[0][2], inter)
    finally:
        lib.hrhPassGeneratorDestroy(pg)


def test_pass_generator_keeps_three_passes_ahead(lib, gpu_backend):
    """Six jobs queued together: the generator renders up to three passes (HRC_MAX_DEFERRED_PASSES) ahead of the readback; every callback
    still sees exactly its own pass count and the sum equals the same passes rendered one after the other."""
    spec = scenes.cornell_box(width=64, height=48, passes=8)
    w, h = spec.width, spec.height
    rg = renderer.Renderer(gpu_backend)
    rg.load_scene(spec)
    for _ in range(6):
        rg.render_pass()
    want = rg.pixels().copy()

    pg = C.c_void_p()
    assert lib.hrhPassGeneratorCreate(0, w, h, C.byref(pg)) == 0, lib.hrhPassGeneratorLastError(pg)
    try:
        def load(user, ctx):
            b = capi.Backend(lib, "hrc", borrowed_ctx=ctx)
            b.width, b.height = w, h
            renderer.Renderer(b).upload_scene(spec, set_environment=False)
        task = capi.CONTEXT_TASK(load)
        lib.hrhPassGeneratorRunTask(pg, task, None)
        got = []

        def done(user, available, pixels, pw, ph, seconds, index):
            got.append((available, index, np.ctypeslib.as_array(pixels, shape=(ph, pw, 4)).copy() if available else None))
        cb = capi.PASS_COMPLETE_CALLBACK(done)
        o = _options(lib, spec, w, h)
        for i in range(6):
            o.resetInternalState = 1 if i == 0 else 0
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert lib.hrhPassGeneratorLastError(pg) == b""
        assert [g[1] for g in got] == [1, 2, 3, 4, 5, 6]
        for k, g in enumerate(got):
            assert g[0] == 1 and (g[2][..., 3] == k + 1).all()
        assert np.allclose(got[-1][2], want, rtol=1e-5, atol=1e-6)
        # a reset in the queue bounds the look-ahead: nothing is rendered ahead across it
        got.clear()
        for i, reset in enumerate([1, 0, 0, 1, 0]):
            o.resetInternalState = reset
            lib.hrhPassGeneratorRenderPass(pg, C.byref(o), cb, None)
        lib.hrhPassGeneratorWaitIdle(pg)
        assert lib.hrhPassGeneratorLastError(pg) == b""
        assert [g[1] for g in got] == [1, 2, 3, 1, 2]
        assert (got[-1][2][..., 3] == 2).all()
    finally:
        lib.hrhPassGeneratorDestroy(pg)
