"""GPU (-m gpu): empty-space skipping.  (1) the occupancy grids (exact per-cell maxima, dilated, octree of maxima) equal
a numpy reduction of the same voxels, whether produced by the lazy reduction kernel (raw uploads, odd
extents, shifted cell origin) or by the decode kernel itself (fused, 256x256x128 bricks); (2) frames and
sample counts are bit-identical with skipping on and off, and match the scalar oracle."""
import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth

pytestmark = pytest.mark.gpu
CS = (32, 32, 16)  # cell size x, y, z


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


def np_grids(a, origin=(0, 0, 0)):
    """raw / dilated / octree maxima of a (Z, Y, X) array with cells aligned to the global grid."""
    Z, Y, X = a.shape
    sh = [origin[i] % CS[i] for i in range(3)]
    n = [(s + h - 1) // c + 1 for s, h, c in zip((X, Y, Z), sh, CS)]
    raw = np.zeros((n[2], n[1], n[0]), np.uint8)
    for cz in range(n[2]):
        for cy in range(n[1]):
            for cx in range(n[0]):
                lo = [max(0, c * s - h) for c, s, h in zip((cx, cy, cz), CS, sh)]
                hi = [min(e, (c + 1) * s - h) for c, s, h, e in zip((cx, cy, cz), CS, sh, (X, Y, Z))]
                blk = a[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]
                raw[cz, cy, cx] = blk.max() if blk.size else 0
    pad = np.pad(raw, ((0, 1), (0, 1), (0, 1)), mode="edge")
    dil = np.zeros_like(raw)
    for dz in range(2):
        for dy in range(2):
            for dx in range(2):
                dil = np.maximum(dil, pad[dz:dz + n[2], dy:dy + n[1], dx:dx + n[0]])
    out = [raw, dil]  # which = 0 raw, 1 = level 0 (dilated), 2.. = octree of maxima above it
    cur = dil
    while cur.size > 1:
        m = [(c - 1) // 2 + 1 for c in cur.shape]
        nxt = np.zeros(m, np.uint8)
        for z in range(m[0]):
            for y in range(m[1]):
                for x in range(m[2]):
                    nxt[z, y, x] = cur[2 * z:2 * z + 2, 2 * y:2 * y + 2, 2 * x:2 * x + 2].max()
        out.append(nxt)
        cur = nxt
    return out


@pytest.mark.parametrize("shape,origin", [((256, 96, 128), (0, 0, 0)), ((70, 50, 100), (0, 0, 0)), ((96, 66, 128), (0, 31, 0)),
                                           ((40, 33, 129), (5, 255, 17))])
def test_lazy_grids_match_numpy(ctx, shape, origin):
    Z, Y, X = shape
    a = synth.rm_region(600, 1000, 870, X, Y, Z)
    v = vr.Volume(ctx, X, Y, Z).set_cell_origin(origin)
    v.upload_brick(a, (0, 0, 0))
    for which, want in enumerate(np_grids(a, origin)):
        assert np.array_equal(v.occupancy(which), want), which
    # overwrite an aligned box and an unaligned one: incremental update resp. lazy rebuild
    b = a.copy()
    if origin == (0, 0, 0) and X >= 64 and Y >= 64 and Z >= 32:
        patch = np.full((16, 32, 32), 201, np.uint8)
        v.upload_brick(patch, (32, 32, 16)); b[16:32, 32:64, 32:64] = 201
        assert np.array_equal(v.occupancy(0), np_grids(b, origin)[0])
    patch = np.full((3, 5, 7), 250, np.uint8)
    v.upload_brick(patch, (1, 2, 3)); b[3:6, 2:7, 1:8] = 250
    for which, want in enumerate(np_grids(b, origin)):
        assert np.array_equal(v.occupancy(which), want), which


def test_decode_keeps_grid_current(ctx):
    """Fused path: full-size bricks, the decode kernel stores each tile's maximum as it writes the tile."""
    bricks = [synth.rm_brick(3, 4, 7), synth.constant_brick(value=0), synth.constant_brick(value=230)]
    streams = [vr.Stream.compress(b, 1, 2)[0].prepare().upload(ctx) for b in bricks]
    vol = vr.Volume(ctx, 512, 256, 256)
    vol.upload_brick(np.zeros((256, 256, 512), np.uint8), (0, 0, 0))
    plan = vr.DecodePlan(ctx, streams, [(0, 0, 0), (256, 0, 0), (256, 0, 128)], vol)
    plan.run()
    dec = vol.download()
    for which, want in enumerate(np_grids(dec)):
        assert np.array_equal(vol.occupancy(which), want), which
    # a second timestep decoded over the first: the grid follows without a rebuild
    s2 = vr.Stream.compress(synth.rm_brick(3, 4, 7, t=2), 1, 2)[0].prepare().upload(ctx)
    plan2 = vr.DecodePlan(ctx, [s2, streams[0]], [(0, 0, 128), (256, 0, 0)], vol)
    plan2.run()
    dec = vol.download()
    for which, want in enumerate(np_grids(dec)):
        assert np.array_equal(vol.occupancy(which), want), which
    # and a frame of it equals the oracle
    p = render.default_params(128, 96, render.COMPOSITE)
    got, S, _ = render.render_to_host(ctx, vol, p)
    want, S2, _ = oa.oracle_render(dec, render.oracle_param_blob(p), 8)
    assert S == S2 and np.abs(got - want).max() < 1e-6


@pytest.mark.parametrize("mode", [render.COMPOSITE, render.ISOSURFACE])
def test_frames_identical_with_and_without_skipping(ctx, mode):
    a = synth.rm_region(600, 1000, 640, 128, 96, 512)  # mostly light fluid (0), layer, heavy fluid (230)
    Z, Y, X = a.shape
    on = vr.Volume(ctx, X, Y, Z); on.upload_brick(a, (0, 0, 0))
    off = vr.Volume(ctx, X, Y, Z).set_skipping(False); off.upload_brick(a, (0, 0, 0))
    views = [render.default_params(192, 144, mode, step=(1 / 128, 1 / 128, 1 / 256)),
             render.oblique_params(192, 144, mode, step=(1 / 256, 1 / 256, 1 / 256)),
             render.default_params(192, 144, mode, eye=(0.1, -0.05, 0.2), front=(0.3, 0.2, 1.0), step=(1 / 100, 1 / 100, 1 / 300)),
             render.default_params(192, 144, mode, eye=(0.05, 0.02, 0.9), front=(0.0, 0.1, -1.0), step=(1 / 256, 1 / 256, 1 / 128))]
    for p in views:
        for iso in ([None] if mode == render.COMPOSITE else [40 / 255, 120 / 255, 228 / 255]):
            if iso is not None:
                p.isoValue = iso
            g1, s1, h1 = render.render_to_host(ctx, on, p, want_hits=True)
            g0, s0, h0 = render.render_to_host(ctx, off, p, want_hits=True)
            assert s1 == s0
            assert np.array_equal(h1, h0)
            assert np.array_equal(g1.view(np.uint32), g0.view(np.uint32))
            want, S, _ = oa.oracle_render(a, render.oracle_param_blob(p), 8)
            assert S == s1
            assert np.nanmax(np.abs(g1 - want)) <= (1e-6 if mode == render.COMPOSITE else 2.0 / 255.0)
