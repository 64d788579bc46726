"""GPU parity (-m gpu): (1) the decode index emitted by the GPU stream producer from its own symbol offsets equals, array
for array, the index the host walk (kd_stream.cpp kd_prepare) derives from the finished stream -- so producer-made streams
never need the walk; (2) random-access decode: any tile subset, from a stream object or
from an indexed container of which only those tiles were read from disk, equals the same voxels of levelCut."""
import glob
import os

import numpy as np
import pytest

import lazy_common as lc
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, render, synth

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]
POISON = 0xEE


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


def _walked(s):
    """The same stream through the host walk only."""
    return vr.Stream.from_memory(s.header, s.distance_map, s.bits).prepare()


def _check_index(a):
    b = _walked(a)
    ia, ib = a.index_arrays(), b.index_arrays()
    for k in ia:
        assert ia[k].shape == ib[k].shape and np.array_equal(ia[k], ib[k]), k
    assert a.index_info() == b.index_info()
    assert np.array_equal(a.serialize_indexed(), b.serialize_indexed())


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_producer_index_on_reference_goldens(ctx, path):
    g = np.load(path)
    a, _ = vr.Stream.compress_device(ctx, g["volume"], int(g["tolerance"]), int(g["maxEpochs"]))
    _check_index(a)


@pytest.mark.parametrize("shape", [(64, 64, 32), (16, 8, 4), (128, 64, 64), (32, 32, 16), (2, 2, 2)])
@pytest.mark.parametrize("tol,epochs", [(1, 2), (6, 1), (0, 5)])
def test_producer_index_small(ctx, shape, tol, epochs):
    X, Y, Z = shape
    for vol in (synth.shell_brick(shape, noise=7, seed=tol + epochs), synth.noise_brick(shape, seed=tol), np.full((Z, Y, X), 9, np.uint8)):
        a, _ = vr.Stream.compress_device(ctx, vol, tol, epochs)
        _check_index(a)


@pytest.mark.parametrize("kind", ["rm", "noise", "constant", "rm_empty"])
def test_producer_index_full_size_brick(ctx, kind):
    vol = {"rm": lambda: synth.rm_brick(3, 4, 7), "noise": synth.noise_brick, "constant": synth.constant_brick,
           "rm_empty": lambda: synth.rm_brick(0, 0, 0)}[kind]()
    a, _ = vr.Stream.compress_device(ctx, vol, 1, 2)
    _check_index(a)
    # and the knob: without the producer's index the stream object carries none until somebody asks
    vr.debug_set("producer_index", 0)
    try:
        b, _ = vr.Stream.compress_device(ctx, vol, 1, 2)
    finally:
        vr.debug_set("producer_index", 1)
    assert np.array_equal(a.bits, b.bits)
    assert np.array_equal(a.serialize_indexed(), b.serialize_indexed())


def _poison(vol, dims):
    X, Y, Z = dims
    vol.upload_brick(np.full((Z, Y, X), POISON, np.uint8), (0, 0, 0))


def _check_subset(got, dec, needed, boxes, stats):
    written = 0
    for b in range(needed.shape[0]):
        for t in range(needed.shape[1]):
            ox, oy, oz = boxes[b, t]
            box = got[oz:oz + 16, oy:oy + 32, ox:ox + 32]
            if needed[b, t]:
                assert np.array_equal(box, dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]), (b, t)
                written += 1
            elif not (box == POISON).all():  # constant member of a needed tile's fill group: written, and correct
                assert np.array_equal(box, dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]), (b, t)
                written += 1
    assert written == stats["tiles_written"]


def test_random_access_decode_of_tile_subsets(ctx, tmp_path):
    dims = (128, 128, 128)
    volume = lc.layered_volume(dims)
    streams, origins = lc.compress_bricks(volume)
    dec = lc.oracle_decode(streams, origins, dims)
    paths = []
    for i, s in enumerate(streams):
        p = tmp_path / f"b{i}.kdx"
        s.save_indexed(p)
        paths.append(p)
    _, boxes = render.lazy_tile_mask(streams, origins, dims, render.default_params(64, 64))
    nt = boxes.shape[1]
    rng = np.random.RandomState(5)
    X, Y, Z = dims
    for trial in range(3):
        needed = (rng.rand(len(streams), nt) < (0.15, 0.5, 0.02)[trial]).astype(np.uint8)
        needed[0, 0] = 1
        # (a) whole streams in memory
        for s in streams:
            s.upload(ctx)
        vol = vr.Volume(ctx, X, Y, Z)
        plan = vr.DecodePlan(ctx, streams, origins, vol)
        _poison(vol, dims)
        st = plan.run_tiles(needed)
        ctx.synchronize()
        with pytest.raises(_lib.VrError) as e:
            vol.download()  # partial
        assert e.value.code == _lib.VR_ERR_STATE
        vr.debug_set("read_partial", 1)
        try:
            got = vol.download()
        finally:
            vr.debug_set("read_partial", 0)
        _check_subset(got, dec, needed, boxes, st)
        plan.run(); ctx.synchronize()
        assert np.array_equal(vol.download(), dec)
        plan.close(); vol.close()
        # (b) the same subset fetched tile by tile from the indexed containers
        subs = [vr.Stream.open_indexed(p, tile_mask=needed[i]) for i, p in enumerate(paths)]
        assert sum(s.bytes_read for s in subs) < sum(os.path.getsize(p) for p in paths)
        for s in subs:
            s.upload(ctx)
        vol = vr.Volume(ctx, X, Y, Z)
        plan = vr.DecodePlan(ctx, subs, origins, vol)
        _poison(vol, dims)
        st2 = plan.run_tiles(needed)
        ctx.synchronize()
        vr.debug_set("read_partial", 1)
        try:
            got2 = vol.download()
        finally:
            vr.debug_set("read_partial", 0)
        _check_subset(got2, dec, needed, boxes, st2)
        assert np.array_equal(got2, got)
        # what is not resident cannot be decoded: a full run and a larger subset are refused
        with pytest.raises(_lib.VrError) as e:
            plan.run()
        assert e.value.code == _lib.VR_ERR_STATE
        more = needed.copy()
        live = [s.index_section("tileLeafBlock") != 0xFFFFFFFF for s in streams]
        cand = [(b, t) for b in range(len(streams)) for t in range(nt) if live[b][t] and not needed[b, t]]
        if cand:
            more[cand[0]] = 1
            with pytest.raises(_lib.VrError) as e:
                plan.run_tiles(more)
            assert e.value.code == _lib.VR_ERR_STATE
        plan.close(); vol.close()


def test_indexed_container_full_open_decodes_like_the_plain_one(ctx, tmp_path):
    vol = synth.rm_brick(3, 4, 7)
    s, _ = vr.Stream.compress_device(ctx, vol, 1, 2)
    p = tmp_path / "rm.kdx"
    s.save_indexed(p)
    t = vr.Stream.open_indexed(p, verify=True)
    out_s = np.empty(vol.size, np.uint8)
    out_t = np.empty(vol.size, np.uint8)
    hd = s.header
    _lib.check(_lib.lib.vr_kd_levelcut(ctx.h, s.h, hd.maxTreeDepth, out_s.ctypes.data))
    _lib.check(_lib.lib.vr_kd_levelcut(ctx.h, t.h, hd.maxTreeDepth, out_t.ctypes.data))
    assert np.array_equal(out_s, out_t)
    assert int(np.abs(out_s.astype(int) - vol.reshape(-1).astype(int)).max()) <= 1  # tolerance 1


def test_view_dependent_frame_from_tiles_fetched_from_disk(ctx, tmp_path):
    """The whole random-access chain (SURVEY 8f-4 + 8f-2): open the containers WITHOUT their streams (index only), ask the index
    which tiles a frame can read (vr_lazy_tile_mask), fetch exactly those tiles' byte ranges from disk, draw the frame with the
    view-dependent decode -- bit-identical to the frame from the fully resident streams."""
    import torch
    dims = (256, 256, 256)
    volume = lc.layered_volume(dims)
    streams, origins = lc.compress_bricks(volume)
    paths = []
    for i, s in enumerate(streams):
        p = tmp_path / f"b{i}.kdx"
        s.save_indexed(p)
        paths.append(p)
    total = sum(os.path.getsize(p) for p in paths)
    X, Y, Z = dims
    for mode, iso in ((render.ISOSURFACE_COMPRESSED, 120 / 255), (render.COMPOSITE, 40 / 255)):
        p = render.oblique_params(192, 160, mode, iso=iso, step=(1 / 256, 1 / 256, 1 / 128))

        def frame_of(ss):
            for s in ss:
                s.upload(ctx)
            vol = vr.Volume(ctx, X, Y, Z)
            plan = vr.DecodePlan(ctx, ss, origins, vol)
            out = torch.zeros((p.height, p.width, 4), dtype=torch.float32, device="cuda:0")
            hit = torch.zeros((p.height, p.width), dtype=torch.int32, device="cuda:0")
            torch.cuda.synchronize()
            st = render.render_compressed_lazy_device(ctx, plan, p, out.data_ptr(), hit.data_ptr(), None)
            ctx.synchronize()
            res = out.cpu().numpy(), hit.cpu().numpy(), st
            plan.close(); vol.close()
            return res

        want_rgba, want_hit, st_full = frame_of(streams)
        nt = streams[0].index_info()["tiles"]
        heads = [vr.Stream.open_indexed(q, tile_mask=np.zeros(nt, np.uint8)) for q in paths]   # header + index, no stream bytes
        head_bytes = sum(h.bytes_read for h in heads)
        needed, _ = render.lazy_tile_mask(heads, origins, dims, p)
        assert 0 < int(needed.sum()) < needed.size
        # the decode keeps whole x-aligned fill groups, whose extra members are constant tiles: they carry no symbols
        subs = [vr.Stream.open_indexed(q, tile_mask=needed[i]) for i, q in enumerate(paths)]
        fetched = sum(s.bytes_read for s in subs)
        assert head_bytes < fetched < total
        got_rgba, got_hit, st = frame_of(subs)
        assert np.array_equal(got_rgba, want_rgba) and np.array_equal(got_hit, want_hit)
        assert st["tiles_written"] == st_full["tiles_written"]
        print(f"\nmode {mode}: {int(needed.sum())} of {needed.size} tiles needed, {fetched} of {total} file bytes fetched")
