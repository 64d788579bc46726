"""GPU parity (-m gpu) of the view-dependent ("lazy") decode behind vr_render_compressed_lazy (the intent
of the reference's stub isosurface_compressed.frag:23-29 and main.cpp:203-237): frames straight from the streams,
decoding only the tiles the frame can read, must be BIT-IDENTICAL to decode-everything-then-march
(vr_render_compressed), while the tiles left out are provably never written (the destination is poisoned first)."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, render, synth

import lazy_common as lc

pytestmark = pytest.mark.gpu
POISON = 0xEE


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def layered(ctx):
    dims = (256, 256, 256)
    vol = lc.layered_volume(dims)
    streams, origins = lc.compress_bricks(vol)
    for s in streams:
        s.upload(ctx)
    return dims, streams, origins, lc.oracle_decode(streams, origins, dims)


def _frame_buffers(p):
    import torch
    out = torch.zeros((p.height, p.width, 4), dtype=torch.float32, device="cuda:0")
    hit = torch.zeros((p.height, p.width), dtype=torch.int32, device="cuda:0")
    smp = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    return out, hit, smp


def _poison(vol, dims):
    X, Y, Z = dims
    vol.upload_brick(np.full((Z, Y, X), POISON, np.uint8), (0, 0, 0))


def _views(mode, iso):
    W, H = 192, 160
    return [render.default_params(W, H, mode, iso=iso, step=(1 / 256, 1 / 256, 1 / 256)),
            render.oblique_params(W, H, mode, iso=iso, step=(1 / 256, 1 / 256, 1 / 128)),
            render.default_params(W, H, mode, iso=iso, eye=(0.2, 0.9, 0.3), front=(-0.1, -1.0, -0.2), up=(0, 0, 1),
                                  step=(1 / 128, 1 / 128, 1 / 200))]


@pytest.mark.parametrize("mode,iso", [(render.ISOSURFACE_COMPRESSED, 40 / 255), (render.ISOSURFACE_COMPRESSED, 120 / 255),
                                      (render.COMPOSITE, 40 / 255)])
def test_lazy_frame_is_bit_identical_and_leaves_culled_tiles_untouched(ctx, layered, mode, iso):
    dims, streams, origins, dec = layered
    X, Y, Z = dims
    vol = vr.Volume(ctx, X, Y, Z)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    for p in _views(mode, iso):
        _poison(vol, dims)
        out, hit, smp = _frame_buffers(p)
        st = render.render_compressed_lazy_device(ctx, plan, p, out.data_ptr(), hit.data_ptr(), smp.data_ptr())
        ctx.synchronize()
        needed, boxes = render.lazy_tile_mask(streams, origins, dims, p)
        assert st["tiles"] == needed.size and 0 < st["tiles_written"] < st["tiles"]
        assert st["tiles_written"] >= int(needed.sum())  # whole groups are kept when one member is needed
        assert st["voxel_bytes"] == st["tiles_written"] * 32 * 32 * 16
        # the volume is partial now: a plain render must refuse it
        with pytest.raises(_lib.VrError) as ei:
            render.render_to_host(ctx, vol, p)
        assert ei.value.code == _lib.VR_ERR_STATE
        with pytest.raises(_lib.VrError) as ei:
            vol.download()
        assert ei.value.code == _lib.VR_ERR_STATE
        vr.debug_set("read_partial", 1)  # look at the stale tiles on purpose
        try:
            got = vol.download()
        finally:
            vr.debug_set("read_partial", 0)
        written = 0
        for b in range(needed.shape[0]):
            for t in range(needed.shape[1]):
                ox, oy, oz = boxes[b, t]
                box = got[oz:oz + 16, oy:oy + 32, ox:ox + 32]
                if needed[b, t]:
                    assert np.array_equal(box, dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]), (b, t)
                    written += 1
                elif (box == POISON).all():
                    pass  # never written
                else:  # member of a kept constant group
                    assert np.array_equal(box, dec[oz:oz + 16, oy:oy + 32, ox:ox + 32]), (b, t)
                    written += 1
        assert written == st["tiles_written"]
        # decode everything, march: the reference semantics of the mode
        out2, hit2, smp2 = _frame_buffers(p)
        render.render_compressed_device(ctx, plan, p, out2.data_ptr(), hit2.data_ptr(), smp2.data_ptr())
        ctx.synchronize()
        assert np.array_equal(vol.download(), dec)
        assert np.array_equal(out.cpu().numpy(), out2.cpu().numpy())
        assert np.array_equal(hit.cpu().numpy(), hit2.cpu().numpy())
        assert int(smp.item()) == int(smp2.item())
        # complete again: the plain march (no bounds from the index, no solid-high mask) is allowed and gives the same frame
        plain, S3, hits3 = render.render_to_host(ctx, vol, p, want_hits=True)
        assert np.array_equal(plain, out2.cpu().numpy()) and np.array_equal(hits3, hit2.cpu().numpy()) and S3 == int(smp2.item())
        if mode != render.COMPOSITE:
            assert (hit2.cpu().numpy() >= 0).any()
    plan.close()
    vol.close()


def test_lazy_matches_the_oracle_frame(ctx, layered):
    """End to end against the CPU restatements: levelCut (plain C) + isosurface.frag restatement."""
    dims, streams, origins, dec = layered
    X, Y, Z = dims
    vol = vr.Volume(ctx, X, Y, Z)
    _poison(vol, dims)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    p = render.oblique_params(192, 160, render.ISOSURFACE_COMPRESSED, iso=40 / 255, step=(1 / 256, 1 / 256, 1 / 128))
    out, hit, smp = _frame_buffers(p)
    render.render_compressed_lazy_device(ctx, plan, p, out.data_ptr(), hit.data_ptr(), smp.data_ptr())
    ctx.synchronize()
    want, S, hits = oa.oracle_render(dec, render.oracle_param_blob(p), 8, want_hits=True)
    assert np.array_equal(hit.cpu().numpy(), hits)
    assert np.abs(out.cpu().numpy() - want).max() <= 2.0 / 255.0
    assert int(smp.item()) == S
    plan.close()
    vol.close()


def test_config3_full_size_64_bricks_1024(ctx):
    """BASELINE.json config 3 at full size: isosurface_compressed straight from the streams of the 64 bricks of layer
    k = 7 (2048 x 2048 x 128) at 1024^2, isoValue 40/255.  Streams come from the GPU producer; a sample of
    bricks is checked bit-exactly against the unmodified reference levelCut (or the plain-C restatement), the lazy
    frame against the full-decode frame (bit-identical) and a sample of rows against the shader restatement."""
    import torch
    BX, BY, BZ = synth.BRICK
    ijk = [(i, j, 7) for j in range(8) for i in range(8)]
    with ThreadPoolExecutor(max_workers=16) as ex:
        bricks = list(ex.map(lambda a: synth.rm_brick(*a), ijk))
    streams, origins = [], []
    for (i, j, _), b in zip(ijk, bricks):
        s, _ = vr.Stream.compress_device(ctx, b, 1, 2)
        s.upload(ctx)
        streams.append(s)
        origins.append((i * BX, j * BY, 0))
    dims = (8 * BX, 8 * BY, BZ)
    vol = vr.Volume(ctx, *dims)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    p = render.default_params(1024, 1024, render.ISOSURFACE_COMPRESSED)  # reference camera, step, isoValue 40/255
    out, hit, smp = _frame_buffers(p)
    st = render.render_compressed_lazy_device(ctx, plan, p, out.data_ptr(), hit.data_ptr(), smp.data_ptr())
    ctx.synchronize()
    assert 0 < st["tiles_written"] < st["tiles"] == 64 * 512
    out2, hit2, smp2 = _frame_buffers(p)
    render.render_compressed_device(ctx, plan, p, out2.data_ptr(), hit2.data_ptr(), smp2.data_ptr())
    ctx.synchronize()
    assert torch.equal(out, out2) and torch.equal(hit, hit2) and int(smp.item()) == int(smp2.item())
    assert int((hit2 >= 0).sum().item()) > 1000
    plain, S3, hits3 = render.render_to_host(ctx, vol, p, want_hits=True)  # plain march of the decoded volume, no index bounds
    assert np.array_equal(plain, out2.cpu().numpy()) and np.array_equal(hits3, hit2.cpu().numpy()) and S3 == int(smp2.item())
    dec = vol.download()
    for b in (0, 27, 63):  # decode parity on a sample of bricks
        s = streams[b]
        hd = s.header
        rc, want = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, BX, BY, BZ)
        assert rc == 0
        ox, oy, oz = origins[b]
        assert np.array_equal(dec[oz:oz + BZ, oy:oy + BY, ox:ox + BX], want)
    # every 64th row of the frame against the shader restatement
    blob = render.oracle_param_blob(p)
    import ctypes
    want = np.zeros((p.height, p.width, 4), np.float32)
    hits = np.full((p.height, p.width), -7, np.int32)
    buf = ctypes.create_string_buffer(blob, len(blob))
    oa.ORACLE.oracle_render_rows(dec.ctypes.data, dims[0], dims[1], dims[2], buf, want.ctypes.data, hits.ctypes.data, 16, 64)
    sel = np.arange(0, p.height, 64)
    got, gh = out2.cpu().numpy(), hit2.cpu().numpy()
    assert np.array_equal(gh[sel], hits[sel])
    assert np.abs(got[sel] - want[sel]).max() <= 2.0 / 255.0
    plan.close()
    vol.close()
