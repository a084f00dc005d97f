"""HRC_OPT_SPECULATIVE_PASSES: hrcRenderPass(i) keeps passes i+1 .. i+n of the same progressive render running ahead in side buffers; a caller
that renders and reads back one pass at a time (the reference viewer's loop, HeatrayRenderer.cpp:366-408) sees the same frames as without
it — exactly its own sample count after every pass, sums equal up to fp32 addition order — whatever it does between two passes."""
import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

pytestmark = pytest.mark.gpu


def _spec():
    return scenes.sphere_grid(width=160, height=90, passes=16, grid=3, u_slices=16, v_slices=8, env_size=(64, 32), max_depth=4)


def _same(a, b):
    assert np.array_equal(a[..., 3], b[..., 3])
    assert np.allclose(a[..., :3], b[..., :3], rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("depth", [1, 3])
def test_speculated_progressive_render_equals_plain(gpu_backend, depth):
    b = gpu_backend
    r = renderer.Renderer(b)
    r.load_scene(_spec())
    plain = []
    for i in range(10):
        r.render_pass(sample_index=i)
        plain.append(r.pixels().copy())
    b.set_option(capi.OPT_SPECULATIVE_PASSES, depth)
    b.clear()
    for i in range(10):
        st = r.render_pass(sample_index=i)
        got = r.pixels()
        assert (got[..., 3] == i + 1).all()                        # pending predictions are invisible: exactly i + 1 samples
        _same(got, plain[i])
        assert st.rays > 0
    with pytest.raises(capi.HrcError):
        b.enqueue_pass_deferred(r.frame)                           # the deferred lanes belong to the speculation
    b.set_option(capi.OPT_SPECULATIVE_PASSES, 0)
    b.clear()
    r.render_pass(sample_index=0)
    _same(r.pixels(), plain[0])                                    # nothing of the discarded predictions leaks in


def test_speculation_stride_follows_a_sharded_rank(gpu_backend):
    """HRC_OPT_SPECULATIVE_STRIDE: rank 1 of 4 renders sample indices 1, 5, 9, 13; the predictions must be those, and be committed."""
    b = gpu_backend
    r = renderer.Renderer(b)
    r.load_scene(_spec())
    idx = [1, 5, 9, 13]
    plain = []
    for i in idx:
        r.render_pass(sample_index=i)
        plain.append(r.pixels().copy())
    with pytest.raises(capi.HrcError):
        b.set_option(capi.OPT_SPECULATIVE_STRIDE, 0)
    b.set_option(capi.OPT_SPECULATIVE_STRIDE, 4)
    b.set_option(capi.OPT_SPECULATIVE_PASSES, 2)
    b.clear(); b.reset_stats()
    for k, i in enumerate(idx):
        r.render_pass(sample_index=i)
        got = r.pixels()
        assert (got[..., 3] == k + 1).all()
        _same(got, plain[k])
    b.set_option(capi.OPT_SPECULATIVE_PASSES, 0)                   # discards what is pending (index 17 is past the 16-entry table: never predicted)
    b.set_option(capi.OPT_SPECULATIVE_STRIDE, 1)
    b.clear()


def test_mispredictions_are_discarded(gpu_backend):
    b = gpu_backend
    r = renderer.Renderer(b)
    r.load_scene(_spec())

    def sequence():
        frames = []
        r.render_pass(sample_index=0); frames.append(r.pixels().copy())
        r.render_pass(sample_index=1); frames.append(r.pixels().copy())
        r.render_pass(sample_index=5); frames.append(r.pixels().copy())        # a jump in the sample index
        g = r.globals; g.maxRayDepth = 2; b.set_globals(g)                     # a state change between two passes
        r.render_pass(sample_index=6); frames.append(r.pixels().copy())
        fov = r.frame.fovTan
        r.frame.fovTan = fov * 1.25                                            # other frame parameters
        r.render_pass(sample_index=7); frames.append(r.pixels().copy())
        b.clear()                                                              # a reset: the render starts over
        r.render_pass(sample_index=0); frames.append(r.pixels().copy())
        r.render_pass(sample_index=15); frames.append(r.pixels().copy())       # last index of the table: nothing left to predict
        r.render_pass(sample_index=15); frames.append(r.pixels().copy())
        r.frame.fovTan = fov
        g.maxRayDepth = 4; b.set_globals(g)
        b.clear()
        return frames

    plain = sequence()
    b.set_option(capi.OPT_SPECULATIVE_PASSES, 3)
    spec = sequence()
    assert len(plain) == len(spec)
    for k, (p, s) in enumerate(zip(plain, spec)):
        bad = ~np.isclose(s[..., :3], p[..., :3], rtol=2e-5, atol=1e-6).all(-1)
        assert not bad.any(), (k, int(bad.sum()), s[bad][:3], p[bad][:3])
        _same(s, p)
