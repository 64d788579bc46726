"""GPU test (-m gpu) of vr_render_to_host_static (csrc/vr_capi_noalloc.cu): the allocation-free variant of
vr_render_to_host, whose device-side frame lives in a static __device__ array of the library.  Checked through the C ABI
against vr_render_to_host on two calls in a row (the array keeps the previous call's contents; the sample counter has to
start at zero each time).  The two entry points run the same march, so the comparison is bit for bit -- tighter than the
1/255 the frames are held to against the oracle (tests/test_gpu_render.py).

Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth

pytestmark = pytest.mark.gpu
TOL = 1.0 / 255.0


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def volume(ctx):
    arr = synth.rm_region(700, 1100, 896, 128, 64, 128)
    Z, Y, X = arr.shape
    v = vr.Volume(ctx, X, Y, Z)
    v.upload_brick(arr, (0, 0, 0))
    yield v
    v.close()


def _same(a, b):
    ga, sa, ha = a
    gb, sb, hb = b
    assert np.abs(ga - gb).max() <= TOL
    assert np.array_equal(ga, gb)
    assert sa == sb and sa > 0
    if ha is not None:
        assert np.array_equal(ha, hb)


def test_static_frame_equals_allocating_frame_two_calls_in_a_row(ctx, volume):
    step = (1 / 96, 1 / 96, 1 / 48)
    # first call: a large composite frame; second call: a smaller isosurface frame over the first one's leftovers
    p1 = render.oblique_params(320, 240, render.COMPOSITE, step=step)
    p2 = render.default_params(160, 120, render.ISOSURFACE, step=step, iso=120 / 255)
    want1 = render.render_to_host(ctx, volume, p1, want_hits=True)
    want2 = render.render_to_host(ctx, volume, p2, want_hits=True)
    got1 = render.render_to_host_static(ctx, volume, p1, want_hits=True)
    got2 = render.render_to_host_static(ctx, volume, p2, want_hits=True)
    _same(got1, want1)
    _same(got2, want2)
    assert (want2[2] >= 0).any()  # the iso frame has hits
    # the same frame again: the sample counter must not carry over
    _same(render.render_to_host_static(ctx, volume, p1, want_hits=True), want1)
    # optional outputs left out
    g, s, h = render.render_to_host_static(ctx, volume, p2, want_hits=False, want_samples=False)
    assert h is None and s == 0 and np.array_equal(g, want2[0])


def test_static_frame_allocates_nothing(ctx, volume):
    import torch
    p = render.default_params(256, 192, render.COMPOSITE, step=(1 / 64, 1 / 64, 1 / 32))
    render.render_to_host_static(ctx, volume, p, want_hits=True)  # module load, symbol address
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        render.render_to_host_static(ctx, volume, p, want_hits=True)
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0


def test_static_frame_rejects_frames_beyond_the_array(ctx, volume):
    n = render.static_max_pixels()
    assert n == 2097152
    p = render.default_params(2048, 1025, render.COMPOSITE)
    assert p.width * p.height > n
    with pytest.raises(vr.VrError) as e:
        render.render_to_host_static(ctx, volume, p)
    assert "VR_STATIC_MAX_PIXELS" in str(e.value)
    # the largest frame that fits goes through (reference window 1600 x 1200 and the full 2 Mi pixels)
    for (w, h) in ((1600, 1200), (2048, 1024)):
        p = render.default_params(w, h, render.COMPOSITE, step=(1 / 32, 1 / 32, 1 / 16))
        _same(render.render_to_host_static(ctx, volume, p, want_hits=True), render.render_to_host(ctx, volume, p, want_hits=True))
