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


def _window_any(grid, below, above):
    """out[k] = any grid[k - below .. k + above] per axis (axes z, y, x of a (cz, cy, cx) array), clipped to the grid."""
    out = grid.copy()
    for axis, (b, a) in zip((2, 1, 0), zip(below, above)):  # x, y, z
        src = out
        acc = np.zeros_like(src)
        n = src.shape[axis]
        for d in range(-b, a + 1):
            lo, hi = max(0, -d), min(n, n - d)
            if lo >= hi:
                continue
            dst_sl = [slice(None)] * 3
            src_sl = [slice(None)] * 3
            dst_sl[axis] = slice(lo, hi)
            src_sl[axis] = slice(lo + d, hi + d)
            acc[tuple(dst_sl)] |= src[tuple(src_sl)]
        out = acc
    return out


@pytest.mark.parametrize("iso,step", [(40 / 255, (1 / 256, 1 / 256, 1 / 256)), (120 / 255, (1 / 64, 1 / 128, 1 / 32)),
                                      (200 / 255, (1 / 512, 1 / 512, 1 / 512))])
def test_iso_rule_equals_a_numpy_restatement(layered, iso, step):
    """The separable window code of kd_cull.h against a direct numpy restatement of the rule:
    fetch = (some cell of c + {0,1}^3 has max > thr) and (some voxel <= thr + 1 within stepReach of c);
    needed = some fetching cell c with the tile in [c - ceil(reach/cs), c + ceil((reach+1)/cs)]."""
    dims, streams, origins, dec = layered
    X, Y, Z = dims
    p = render.default_params(64, 64, render.ISOSURFACE, iso=iso, step=step)
    needed, boxes = render.lazy_tile_mask(streams, origins, dims, p)
    thr = int(np.ceil(iso * 255.0)) - 1
    cs = (32, 32, 16)
    G = (X, Y, Z)
    cmax = dec.reshape(Z // 16, 16, Y // 32, 32, X // 32, 32).max(axis=(1, 3, 5))
    cmin = dec.reshape(Z // 16, 16, Y // 32, 32, X // 32, 32).min(axis=(1, 3, 5))
    ceil_div = lambda a, b: -(-a // b)
    reach = [int(np.ceil((abs(s) + 0.01) * g)) + 1 for s, g in zip(step, G)]
    sreach = [int(np.ceil(abs(s) * g)) + 1 for s, g in zip(step, G)]
    hot, low = cmax > thr, cmin <= thr + 1
    fetch = _window_any(hot, (0, 0, 0), (1, 1, 1))
    fetch &= _window_any(low, [ceil_div(r, c) for r, c in zip(sreach, cs)], [ceil_div(r + 1, c) for r, c in zip(sreach, cs)])
    want = _window_any(fetch, [ceil_div(r + 1, c) for r, c in zip(reach, cs)], [ceil_div(r, c) for r, c in zip(reach, cs)])
    got = np.zeros_like(want)
    for b in range(needed.shape[0]):
        for t in range(needed.shape[1]):
            cx, cy, cz = boxes[b, t] // np.array(cs)
            got[cz, cy, cx] = bool(needed[b, t])
    assert np.array_equal(got, want)
    assert 0 < got.sum() < got.size


@pytest.mark.parametrize("kernel", ["persistent", "classic"])
def test_culled_work_list_on_the_cpu_emulation(kernel):
    """The device path's work-list builders with a cull mask (kd_worklist.h), run by the CPU emulation of the decode
    kernel (tests/emu): tiles the mask keeps are decoded bit-exactly, every member of a kept constant group too, all
    other tiles are never written, and their (cell, exact maximum) pairs are reported for the occupancy grid."""
    import ctypes
    if oa.EMU is None or not hasattr(oa.EMU, "emu_set_lazy"):
        pytest.skip("emulation harness not built")
    X, Y, Z = 256, 256, 128
    brick = synth.rm_brick(3, 4, 7)  # straddles the mixing layer: void, mixed and solid tiles
    s, _ = vr.Stream.compress(brick, 1, 2)
    p = render.default_params(64, 64, render.ISOSURFACE, iso=40 / 255)
    needed, boxes = render.lazy_tile_mask([s], [(0, 0, 0)], (X, Y, Z), p)
    needed = np.ascontiguousarray(needed[0])
    assert 0 < needed.sum() < needed.size
    mins, maxs, _ = render.stream_tile_bounds(s)
    hd, dm, bits = s.header, s.distance_map, s.bits
    rc, want = oa.oracle_levelcut(bits, hd.numActiveNodes, dm, hd.origTreeDepth, hd.maxTreeDepth, X, Y, Z)
    assert rc == 0
    h = oa.EMU.emu_from_memory(ctypes.byref(hd), dm.ctypes.data, bits.ctypes.data, bits.size)
    slab = np.full((Z, Y, X), 0xEE, np.uint8)
    culled = np.zeros((needed.size, 2), np.int64)
    oa.EMU.emu_set_lazy(needed.ctypes.data, culled.ctypes.data, needed.size)
    try:
        if kernel == "persistent":
            assert oa.EMU.emu_decode6(h, slab.ctypes.data, X, Y, Z, 0, 0, 0, 304) == 0
        else:
            assert oa.EMU.emu_decode(h, slab.ctypes.data, X, Y, Z, 0, 0, 0, 1) == 0
        n = oa.EMU.emu_culled_count()
    finally:
        oa.EMU.emu_set_lazy(None, None, 0)
        oa.EMU.emu_free(h)
    cells = {}
    for t, (ox, oy, oz) in enumerate(boxes[0]):
        cells[(ox >> 5) + 8 * ((oy >> 5) + 8 * (oz >> 4))] = t
    dropped = set()
    for cell, value in culled[:n]:
        t = cells[int(cell)]
        assert not needed[t] and int(value) == int(maxs[t])
        dropped.add(t)
    assert len(dropped) == n > 0
    for t, (ox, oy, oz) in enumerate(boxes[0]):
        got = slab[oz:oz + 16, oy:oy + 32, ox:ox + 32]
        if t in dropped:
            assert (got == 0xEE).all(), t
        else:  # needed, or a member of a constant group that has a needed member
            assert np.array_equal(got, want[oz:oz + 16, oy:oy + 32, ox:ox + 32]), t
    assert not any(needed[t] for t in dropped)
