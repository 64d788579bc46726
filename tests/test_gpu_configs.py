"""GPU parity (-m gpu) at the shapes BASELINE.json names (configs 2, 4, 5; config 3 lives in test_gpu_lazy.py), on ONE
device.  Decode is checked bit-exactly against the levelCut restatement (pinned to the unmodified reference by
tests/test_oracle.py) on whole bricks, frames against the scalar shader restatement on a sample of rows with the
tolerances of SURVEY B.5 (RGBA max-abs <= 1/255), and at full size through properties that need no CPU pass over
7.5 GiB: the slab-sharded decode equals the single-texture decode byte for byte, and the sort-last composite of 8
virtual ranks equals the single-texture frame."""
import ctypes
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, sortlast, synth

pytestmark = pytest.mark.gpu
TOL = 1.0 / 255.0
BX, BY, BZ = synth.BRICK


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0, cuda_stream=torch.cuda.current_stream().cuda_stream)
    yield c
    c.close()


def _levelcut(s):
    hd = s.header
    rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, hd.X, hd.Y, hd.Z)
    assert rc == 0
    return dec


def _oracle_rows(vol_host, p, row_step, want_hits=False):
    """Shader restatement on rows py % row_step == 0 only (oracle_render_rows)."""
    Z, Y, X = vol_host.shape
    blob = render.oracle_param_blob(p)
    buf = ctypes.create_string_buffer(blob, len(blob))
    out = np.zeros((p.height, p.width, 4), np.float32)
    hits = np.zeros((p.height, p.width), np.int32)
    oa.ORACLE.oracle_render_rows(vol_host.ctypes.data, X, Y, Z, buf, out.ctypes.data, hits.ctypes.data, 16, row_step)
    sel = np.arange(0, p.height, row_step)
    return sel, out[sel], hits[sel]


def _torch_volume(ctx, dims):
    """A device slab that torch can look at: the library wraps torch's allocation (vr_volume_wrap)."""
    X, Y, Z = dims
    t = torch.zeros((Z, Y, X), dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    return t, vr.Volume(ctx, X, Y, Z, device_ptr=t.data_ptr())


def test_config2_single_brick_decode_and_composite_512(ctx):
    """Config 2: RM brick (3,4,7) of the synthetic field (b = 483, straddles the interface), GPU decode + colour
    compositing at 512^2 against levelCut + raycaster.frag restated on the CPU."""
    brick = synth.rm_brick(3, 4, 7)
    s, _ = vr.Stream.compress_device(ctx, brick, 1, 2)
    s.upload(ctx)
    vol = vr.Volume(ctx, BX, BY, BZ)
    plan = vr.DecodePlan(ctx, [s], [(0, 0, 0)], vol)
    plan.run()
    got = vol.download()
    want = _levelcut(s)
    assert np.array_equal(got, want)
    if oa.REF is not None:  # the unmodified reference decodes its own build of the same brick to the same voxels
        ref = oa.ref_build(brick, 1, 2)
        assert ref["n"] == s.header.numActiveNodes and np.array_equal(oa.ref_levelcut(ref), want)
    for p in (render.default_params(512, 512, render.COMPOSITE), render.oblique_params(512, 512, render.COMPOSITE)):
        frame, S, _ = render.render_to_host(ctx, vol, p)
        ref_frame, S2, _ = oa.oracle_render(want, render.oracle_param_blob(p), 16)
        assert S == S2
        assert np.abs(frame - ref_frame).max() <= TOL
    plan.close(); vol.close()


def _build_streams(ctx, ijk, t=0, tol=1, epochs=2):
    """GPU producer for the mixed bricks, one shared stream per constant value; prepared on host threads."""
    const = {}
    mixed = [a for a in ijk if synth.rm_brick_constant_value(*a, t) is None]
    with ThreadPoolExecutor(max_workers=16) as ex:
        vols = dict(zip(mixed, ex.map(lambda a: synth.rm_brick(*a, t), mixed)))
    out = {}
    for a in ijk:
        cv = synth.rm_brick_constant_value(*a, t)
        if cv is None:
            out[a], _ = vr.Stream.compress_device(ctx, vols[a], tol, epochs)
        else:
            if cv not in const:
                const[cv], _ = vr.Stream.compress(synth.constant_brick(value=cv), tol, epochs)
            out[a] = const[cv]
    uniq = list({id(s): s for s in out.values()}.values())
    with ThreadPoolExecutor(max_workers=16) as ex:
        list(ex.map(lambda s: s.prepare(), uniq))
    for s in uniq:
        s.upload(ctx)
    return out, vols


@pytest.fixture(scope="module")
def cfg4(ctx):
    """Config 4 at full size, shared by the composite and the isosurface test: 2048 x 2048 x 1920, 960 bricks, tolerance
    1 / 2 epochs, decoded once as a single texture and once brick-sharded into 8 y-slabs (one per virtual rank) whose
    halo is wide enough for both modes (composite 1 row, isosurface 31 rows at 2048: SURVEY H7)."""
    GI, GJ, GK = synth.GRID
    ijk = [(i, j, k) for k in range(GK) for j in range(GJ) for i in range(GI)]
    streams, vols = _build_streams(ctx, ijk)
    G = (GI * BX, GJ * BY, GK * BZ)
    full_t, full = _torch_volume(ctx, G)
    plan = vr.DecodePlan(ctx, [streams[a] for a in ijk], [(i * BX, j * BY, k * BZ) for (i, j, k) in ijk], full)
    plan.run()
    ctx.synchronize()
    W = H = 1024
    views = {m: [render.default_params(W, H, m), render.oblique_params(W, H, m)] for m in (render.COMPOSITE, render.ISOSURFACE)}
    n = 8
    halos = [sortlast.required_halo(p, G) for m in views for p in views[m]]
    halo = [max(h[a] for h in halos) for a in range(3)]
    assert halo[1] >= 31  # the isosurface's look-ahead sample + gradient probes at 2048 (isosurface.frag:18,52-59,121)
    descs, stored = sortlast.slab_layout(G, n, axis=1, halo=halo)
    ops, keep = [], []
    for r in range(n):
        o, e = stored[r]
        slab_t, slab = _torch_volume(ctx, e)
        slab.set_cell_origin(o)
        mine = [a for a in ijk if a[1] == r]
        pl = vr.DecodePlan(ctx, [streams[a] for a in mine], [(i * BX - o[0], j * BY - o[1], k * BZ - o[2]) for (i, j, k) in mine], slab)
        pl.run()
        keep.append((slab_t, pl))
        ops.append(sortlast.CudaSlabOps(ctx, sortlast.SortLast(ctx, descs, r), slab, W, H))
    sortlast.exchange_halos_lockstep(ops, descs, stored)
    ctx.synchronize(); torch.cuda.synchronize()
    host = {}

    def host_volume():  # one 7.5 GiB download, shared by the row checks of both tests
        if "v" not in host:
            host["v"] = full_t.cpu().numpy()
        return host["v"]

    yield dict(ijk=ijk, streams=streams, G=G, full_t=full_t, full=full, views=views, descs=descs, stored=stored, ops=ops, keep=keep,
               W=W, H=H, host_volume=host_volume)
    for _, pl in keep:
        pl.close()
    plan.close()


def test_config4_full_volume_960_bricks_sharded_over_8_virtual_ranks(ctx, cfg4):
    """Config 4, COMPOSITE.  Properties: slabs == single texture byte for byte; a sample of bricks bit-exact vs levelCut;
    sort-last frame == single frame (fp32 re-association, < 1/255); rows of the single frame vs the shader restatement on
    the downloaded volume."""
    c = cfg4
    full_t, full, streams = c["full_t"], c["full"], c["streams"]
    for a in [(0, 0, 0), (3, 4, 7), (7, 7, 7), (5, 2, 8), (2, 6, 14)]:  # constant below, mixing layer, constant above
        i, j, k = a
        want = torch.from_numpy(_levelcut(streams[a])).cuda()
        assert torch.equal(full_t[k * BZ:(k + 1) * BZ, j * BY:(j + 1) * BY, i * BX:(i + 1) * BX], want), a
    W, H = c["W"], c["H"]
    views = c["views"][render.COMPOSITE]
    single = []
    for p in views:
        out = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
        torch.cuda.synchronize()
        render.render_device(ctx, full, p, out.data_ptr())
        ctx.synchronize()
        single.append(out)
    for r in range(8):  # owned rows + halo rows == the same rows of the single texture
        o, e = c["stored"][r]
        assert torch.equal(c["keep"][r][0], full_t[o[2]:o[2] + e[2], o[1]:o[1] + e[1], o[0]:o[0] + e[0]]), r
    for p, one in zip(views, single):
        frames = sortlast.render_frame_lockstep(c["ops"], p)
        torch.cuda.synchronize()
        for f in frames:
            assert float((f - one).abs().max().item()) < 1e-4 < TOL
    # the single-texture frame against the shader restatement, every 64th row, on the decoded volume
    host = c["host_volume"]()
    for p, one in zip(views, single):
        sel, want, _ = _oracle_rows(host, p, 64)
        assert np.abs(one.cpu().numpy()[sel] - want).max() <= TOL


def test_config4_full_volume_isosurface_sharded_over_8_virtual_ranks(ctx, cfg4):
    """Config 4, ISOSURFACE ("COMPOSITE (and ISOSURFACE)") at full size with the 31-row halo of SURVEY H7: the
    sort-last first-hit frame of 8 virtual ranks is bit-identical to the single-texture frame (same taps, same
    arithmetic: isosurface.frag:120-157), for the reference camera and the oblique view whose rays cross the slabs; rows
    of the single frame against the shader restatement (hit / miss classification equal, hit index within one step,
    colour <= 2/255 on agreeing hits: SURVEY B.5)."""
    c = cfg4
    W, H, full = c["W"], c["H"], c["full"]
    host = c["host_volume"]()
    for p in c["views"][render.ISOSURFACE]:
        one = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
        hit = torch.zeros((H, W), dtype=torch.int32, device="cuda:0")
        torch.cuda.synchronize()
        render.render_device(ctx, full, p, one.data_ptr(), hit.data_ptr())
        ctx.synchronize()
        assert int((hit >= 0).sum().item()) > 1000  # the interface is in view
        frames = sortlast.render_frame_lockstep(c["ops"], p)
        torch.cuda.synchronize()
        for r, f in enumerate(frames):
            same = (f == one) | (torch.isnan(f) & torch.isnan(one))
            assert bool(same.all().item()), (r, int((~same).sum().item()))
        sel, ref, ref_hits = _oracle_rows(host, p, 64)
        got, gh = one.cpu().numpy()[sel], hit.cpu().numpy()[sel]
        assert np.array_equal(gh >= 0, ref_hits >= 0) and (gh >= 0).any()
        assert np.abs(gh - ref_hits).max() <= 1
        assert np.nanmax(np.abs(got - ref)[gh == ref_hits]) <= 2 * TOL


SWEEP = [(1, 6), (2, 4), (2, 2), (2, 1), (5, 1), (5, 0)]  # (maxEpochs, tolerance), SURVEY 8d config 5


def test_config5_timestep_sweep_2048(ctx):
    """Config 5 on one device, bounded: timesteps t = 0..7 of the brick column (3, 4, k = 4..10) -- the bricks the
    growing mixing layer crosses -- each timestep with the next (maxEpochs, tolerance) pair of the sweep, decoded and
    drawn at 2048^2.  Per timestep: every brick bit-exact vs levelCut; rows of the frame vs the shader restatement."""
    ks = list(range(4, 11))
    dims = (BX, BY, BZ * len(ks))
    vol_t, vol = _torch_volume(ctx, dims)
    out = torch.zeros((2048, 2048, 4), dtype=torch.float32, device="cuda:0")
    for t in range(8):
        epochs, tol = SWEEP[t % len(SWEEP)]
        ijk = [(3, 4, k) for k in ks]
        streams, _ = _build_streams(ctx, ijk, t=t, tol=tol, epochs=epochs)
        plan = vr.DecodePlan(ctx, [streams[a] for a in ijk], [(0, 0, n * BZ) for n in range(len(ks))], vol)
        plan.run()
        ctx.synchronize()
        with ThreadPoolExecutor(max_workers=8) as ex:
            want = list(ex.map(lambda a: _levelcut(streams[a]), ijk))
        host = np.concatenate(want, axis=0)
        assert torch.equal(vol_t, torch.from_numpy(host).cuda()), t
        p = render.oblique_params(2048, 2048, render.COMPOSITE if t % 2 == 0 else render.ISOSURFACE,
                                  step=(1 / 256, 1 / 256, 1 / (128 * len(ks))))
        hit = torch.zeros((2048, 2048), dtype=torch.int32, device="cuda:0")
        torch.cuda.synchronize()
        render.render_device(ctx, vol, p, out.data_ptr(), hit.data_ptr())
        ctx.synchronize()
        sel, ref, ref_hits = _oracle_rows(host, p, 128)
        got = out.cpu().numpy()[sel]
        if p.mode == render.COMPOSITE:
            assert np.abs(got - ref).max() <= TOL, t
        else:
            gh = hit.cpu().numpy()[sel]
            assert np.array_equal(gh >= 0, ref_hits >= 0) and (gh >= 0).any(), t
            assert np.abs(gh - ref_hits).max() <= 1, t
            assert np.abs(got - ref)[gh == ref_hits].max() <= 2 * TOL, t
        plan.close()
