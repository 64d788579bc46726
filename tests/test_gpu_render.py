"""GPU parity (-m gpu): CUDA ray-marching through the C ABI vs the scalar shader restatement
(oracle/render_oracle.cpp).  Stated tolerances: composite RGBA max-abs <= 1/255; iso
hit/miss identical and hit index within +-1 step, colour <= 2/255 on agreeing hits.  The kernel
keeps the oracle's fp32 operation order, so the expected difference is ~0 (powf aside)."""
import numpy as np
import pytest

import oracle_api as oa
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
def volumes(ctx):
    out = {}
    for name, arr in (("shell", synth.shell_brick((64, 64, 32), noise=5)),
                      ("rm", synth.rm_region(700, 1100, 896, 128, 64, 128)),
                      ("tiny", synth.noise_brick((4, 2, 8), seed=5))):
        Z, Y, X = arr.shape
        v = vr.Volume(ctx, X, Y, Z)
        v.upload_brick(arr, (0, 0, 0))
        out[name] = (arr, v)
    return out


def _views(W, H, mode):
    return [render.default_params(W, H, mode, step=(1 / 64, 1 / 64, 1 / 32)),
            render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 48)),
            render.default_params(W, H, mode, eye=(0.1, -0.05, 0.2), front=(0.3, 0.2, 1.0)),  # eye inside the cube
            render.default_params(W, H, mode, eye=(0, 0, -0.58), step=(1 / 256, 1 / 256, 1 / 128))]  # front face near-clipped


@pytest.mark.parametrize("name", ["shell", "rm", "tiny"])
def test_composite_matches_oracle(ctx, volumes, name):
    arr, v = volumes[name]
    for p in _views(160, 120, render.COMPOSITE):
        got, S, _ = render.render_to_host(ctx, v, p)
        want, S2, _ = oa.oracle_render(arr, render.oracle_param_blob(p), 8)
        assert S == S2
        assert np.abs(got - want).max() <= TOL
        assert np.array_equal(np.round(got * 255), np.round(want * 255)) or np.abs(got - want).max() < 1e-5


@pytest.mark.parametrize("name", ["shell", "rm"])
@pytest.mark.parametrize("iso", [40 / 255, 120 / 255])
def test_isosurface_matches_oracle(ctx, volumes, name, iso):
    arr, v = volumes[name]
    for p in _views(160, 120, render.ISOSURFACE):
        p.isoValue = iso
        got, S, hits = render.render_to_host(ctx, v, p, want_hits=True)
        want, S2, hits2 = oa.oracle_render(arr, render.oracle_param_blob(p), 8, want_hits=True)
        assert np.array_equal(hits >= 0, hits2 >= 0)
        assert np.abs(hits - hits2).max() <= 1
        agree = hits == hits2
        assert np.abs(got - want)[agree].max() <= 2 * TOL
        assert abs(S - S2) <= (~agree).sum()


def test_render_compressed_is_iso_of_decoded_streams(ctx):
    """ISOSURFACE_COMPRESSED semantics: isosurface.frag over levelCut(streams), all on device."""
    bx, by, bz = 32, 32, 16
    streams, origins, host = [], [], np.zeros((bz, 2 * by, 2 * bx), np.uint8)
    for b in range(4):
        i, j = b % 2, b // 2
        brick = synth.rm_region(640 + i * bx, 900 + j * by, 950, bx, by, bz)
        s, _ = vr.Stream.compress(brick, 1, 2)
        s.upload(ctx)
        streams.append(s); origins.append((i * bx, j * by, 0))
        hd = s.header
        rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, bx, by, bz)
        assert rc == 0
        host[:, j * by:(j + 1) * by, i * bx:(i + 1) * bx] = dec
    vol = vr.Volume(ctx, 2 * bx, 2 * by, bz)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    p = render.oblique_params(128, 96, render.ISOSURFACE_COMPRESSED, step=(1 / 64, 1 / 64, 1 / 16), iso=100 / 255)
    import torch
    out = torch.empty((96, 128, 4), dtype=torch.float32, device="cuda:0")
    hit = torch.empty((96, 128), dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    render.render_compressed_device(ctx, plan, p, out.data_ptr(), hit.data_ptr())
    ctx.synchronize()
    want, _, hits2 = oa.oracle_render(host, render.oracle_param_blob(p), 8, want_hits=True)
    assert np.array_equal(hit.cpu().numpy(), hits2)
    assert np.abs(out.cpu().numpy() - want).max() <= 2 * TOL
    assert (hits2 >= 0).any()
