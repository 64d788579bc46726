"""CPU checks of the view-dependent ("lazy") decode decision (csrc/kd_cull.h): the per-tile bounds the
stream index carries are exact, and the tiles the rule leaves out cannot influence a frame -- checked with the scalar
shader restatement (oracle/render_oracle.cpp), which fetches at every lattice position: overwriting every culled tile
with other values below the iso value must leave frame, hit indices and sample count bit-identical."""
import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth

import lazy_common as lc


@pytest.mark.parametrize("shape,kind", [((64, 64, 32), "shell"), ((128, 128, 64), "rm"), ((64, 64, 32), "noise"),
                                        ((64, 64, 32), "constant"), ((256, 256, 128), "rm")])
def test_tile_bounds_are_exact(shape, kind):
    X, Y, Z = shape
    if kind == "shell":
        brick = synth.shell_brick(shape, noise=4)
    elif kind == "rm":
        brick = synth.rm_region(768, 1024, 896, X, Y, Z)
    elif kind == "noise":
        brick = synth.noise_brick(shape, seed=3)
    else:
        brick = synth.constant_brick(shape, 91)
    s, _ = vr.Stream.compress(brick, 1, 2)
    mins, maxs, td = render.stream_tile_bounds(s)
    assert td == (32, 32, 16)
    hd = s.header
    rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, X, Y, Z)
    assert rc == 0
    p = render.default_params(16, 16, render.ISOSURFACE)
    _, boxes = render.lazy_tile_mask([s], [(0, 0, 0)], shape, p)
    assert boxes.shape == (1, mins.size, 3)
    seen = np.zeros((Z, Y, X), bool)
    for t, (ox, oy, oz) in enumerate(boxes[0]):
        box = dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]
        assert box.min() == mins[t] and box.max() == maxs[t], t
        seen[oz:oz + 16, oy:oy + 32, ox:ox + 32] = True
    assert seen.all()


@pytest.fixture(scope="module")
def layered():
    dims = (256, 256, 256)
    vol = lc.layered_volume(dims)
    streams, origins = lc.compress_bricks(vol)
    return dims, streams, origins, lc.oracle_decode(streams, origins, dims)


def _views(mode, iso):
    W, H = 160, 128
    return [render.default_params(W, H, mode, iso=iso, step=(1 / 256, 1 / 256, 1 / 256)),
            render.oblique_params(W, H, mode, iso=iso, step=(1 / 256, 1 / 256, 1 / 128)),
            render.default_params(W, H, mode, iso=iso, eye=(0.2, 0.9, 0.3), front=(-0.1, -1.0, -0.2), up=(0, 0, 1),
                                  step=(1 / 128, 1 / 128, 1 / 200))]


@pytest.mark.parametrize("iso", [40 / 255, 120 / 255, 200 / 255])
def test_culled_tiles_cannot_influence_an_iso_frame(layered, iso):
    dims, streams, origins, dec = layered
    rng = np.random.default_rng(5)
    for p in _views(render.ISOSURFACE, iso):
        needed, boxes = render.lazy_tile_mask(streams, origins, dims, p)
        thr = int(np.ceil(iso * 255.0)) - 1  # largest voxel value strictly below the iso value
        assert 0 < needed.sum() < needed.size  # something is culled, something is kept
        poisoned = dec.copy()
        kinds = set()
        for b in range(needed.shape[0]):
            for t in np.flatnonzero(needed[b] == 0):
                ox, oy, oz = boxes[b, t]
                box = dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]
                if box.max() <= thr:     # void: below the iso value whatever is stored there
                    poisoned[oz:oz + 16, oy:oy + 32, ox:ox + 32] = rng.integers(0, thr + 1, (16, 32, 32), dtype=np.uint8)
                    kinds.add("void")
                else:                    # solid-high: at or above it whatever is stored there
                    assert box.min() >= thr + 2
                    poisoned[oz:oz + 16, oy:oy + 32, ox:ox + 32] = rng.integers(thr + 2, 256, (16, 32, 32), dtype=np.uint8)
                    kinds.add("solid")
        assert (poisoned != dec).any()
        assert kinds == {"void", "solid"}
        want, S, hits = oa.oracle_render(dec, render.oracle_param_blob(p), 8, want_hits=True)
        got, S2, hits2 = oa.oracle_render(poisoned, render.oracle_param_blob(p), 8, want_hits=True)
        assert (hits >= 0).any()
        assert S == S2 and np.array_equal(hits, hits2) and np.array_equal(want, got)


def test_composite_keeps_every_tile_next_to_a_nonzero_one(layered):
    """Composite frames skip exactly-zero cells only: a tile is left out iff it and its 26 neighbours (+ one more cell
    towards -x,-y,-z, the dilation of the occupancy pyramid) hold nothing but zeros."""
    dims, streams, origins, dec = layered
    p = render.default_params(64, 64, render.COMPOSITE)
    needed, boxes = render.lazy_tile_mask(streams, origins, dims, p)
    assert 0 < needed.sum() < needed.size
    X, Y, Z = dims
    cellmax = dec.reshape(Z // 16, 16, Y // 32, 32, X // 32, 32).max(axis=(1, 3, 5))  # (cz, cy, cx)
    for b in range(needed.shape[0]):
        for t in range(needed.shape[1]):
            cx, cy, cz = boxes[b, t] // np.array([32, 32, 16])
            # hot cell H marks [H - 1, H + 1] for composite, so tile c is needed iff some cell in [c - 1, c + 1] is nonzero
            nb = cellmax[max(cz - 1, 0):cz + 2, max(cy - 1, 0):cy + 2, max(cx - 1, 0):cx + 2]
            assert bool(needed[b, t]) == bool(nb.max() > 0), (b, t)
