"""GPU parity (-m gpu) of the sort-last renderer: N virtual ranks on ONE device run the product's slab
kernels through the C ABI in lock-step (the collectives done locally), and the composited frame is
compared with (a) the scalar single-texture oracle of the whole volume and (b) the product's own
single-GPU frame.  Tolerances: composite RGBA max-abs <= 1/255 vs the oracle (expected ~1e-6: the
segment algebra reorders fp32 products); isosurface identical to the single-GPU frame and within the
single-GPU tolerances (hit mask identical, colour <= 2/255) of the oracle."""
import numpy as np
import pytest
import torch

import oracle_api as oa
import volumerenderer_b200 as vr
from sortlast_cpu import cut_slab
from volumerenderer_b200 import render, sortlast, synth

pytestmark = pytest.mark.gpu
TOL = 1.0 / 255.0
W, H = 160, 120


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0, cuda_stream=torch.cuda.current_stream().cuda_stream)  # same stream as torch's ops
    yield c
    c.close()


@pytest.fixture(scope="module")
def full():
    return synth.rm_region(640, 1000, 900, 64, 64, 96)


def _views(mode):
    return [render.default_params(W, H, mode, step=(1 / 64, 1 / 64, 1 / 96)),
            render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 128)),
            render.default_params(W, H, mode, eye=(0.2, 0.9, -0.4), front=(-0.2, -0.9, 0.4), up=(0, 0, 1), step=(1 / 64, 1 / 64, 1 / 96)),
            render.default_params(W, H, mode, eye=(0.1, -0.05, 0.2), front=(0.3, 0.2, 1.0), step=(1 / 64, 1 / 64, 1 / 96))]  # eye inside


def _make_ranks(ctx, full, n, axis, p, fill_halo_by_exchange=True):
    Z, Y, X = full.shape
    halo = sortlast.required_halo(p, (X, Y, Z))
    descs, stored = sortlast.slab_layout((X, Y, Z), n, axis=axis, halo=halo)
    ops = []
    for r in range(n):
        o, e = stored[r]
        vol = vr.Volume(ctx, *e).set_cell_origin(o)
        lo = [descs[r].own_lo[a] for a in range(3)]; hi = [descs[r].own_hi[a] for a in range(3)]
        own = np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]])
        torch.cuda.synchronize()
        vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))
        vol.upload_brick(own, [lo[a] - o[a] for a in range(3)])
        ops.append(sortlast.CudaSlabOps(ctx, sortlast.SortLast(ctx, descs, r), vol, W, H, count_samples=True))
    sortlast.exchange_halos_lockstep(ops, descs, stored)
    torch.cuda.synchronize()
    for r in range(n):
        assert np.array_equal(ops[r].vol.download(), cut_slab(full, descs[r], stored[r])), "halo exchange"
    return ops


@pytest.mark.parametrize("n,axis", [(2, 1), (4, 1), (8, 1), (3, 2), (2, 0)])
def test_composite_sortlast(ctx, full, n, axis):
    Z, Y, X = full.shape
    single = vr.Volume(ctx, X, Y, Z)
    single.upload_brick(full, (0, 0, 0))
    for p in _views(render.COMPOSITE):
        ops = _make_ranks(ctx, full, n, axis, p)
        frames = sortlast.render_frame_lockstep(ops, p)
        torch.cuda.synchronize()
        want, S, _ = oa.oracle_render(full, render.oracle_param_blob(p), 8)
        one, S1, _ = render.render_to_host(ctx, single, p)
        for f in frames:
            got = f.cpu().numpy()
            assert np.abs(got - want).max() <= TOL
            assert np.abs(got - want).max() < 2e-5
            assert np.abs(got - one).max() < 2e-5
        # every lattice sample is fetched by exactly one rank in pass 1 unless its slab stopped early
        # (segment alone opaque); pass 2 re-marches at most one segment per ray
        n1 = sum(int(o.samples.item()) for o in ops)
        assert n1 >= S1 and n1 <= 3 * S


@pytest.mark.parametrize("n,axis", [(2, 1), (4, 1), (3, 2)])
@pytest.mark.parametrize("iso", [40 / 255, 120 / 255])
def test_isosurface_sortlast(ctx, full, n, axis, iso):
    Z, Y, X = full.shape
    single = vr.Volume(ctx, X, Y, Z)
    single.upload_brick(full, (0, 0, 0))
    for p in _views(render.ISOSURFACE):
        p.isoValue = iso
        ops = _make_ranks(ctx, full, n, axis, p)
        frames = sortlast.render_frame_lockstep(ops, p)
        torch.cuda.synchronize()
        one, _, hits1 = render.render_to_host(ctx, single, p, want_hits=True)
        want, _, hits = oa.oracle_render(full, render.oracle_param_blob(p), 8, want_hits=True)
        assert np.array_equal(hits1 >= 0, hits >= 0)
        for f in frames:
            got = f.cpu().numpy()
            assert np.array_equal(np.isnan(got), np.isnan(one))
            assert np.array_equal(np.nan_to_num(got), np.nan_to_num(one))  # same kernel arithmetic, same taps
            assert np.nanmax(np.abs(got - want)) <= 2 * TOL


def test_compressed_bricks_decode_into_slabs(ctx):
    """Config-4 shape in miniature: 2x2x2 bricks of 32^3, compressed, sharded by brick row over 2 ranks;
    each rank decodes its bricks in place inside its halo'd slab, halos travel, frame is composited."""
    full = synth.rm_region(700, 1100, 930, 64, 64, 64)
    B = 32
    p = render.oblique_params(W, H, render.COMPOSITE, step=(1 / 96, 1 / 96, 1 / 96))
    descs, stored = sortlast.slab_layout((64, 64, 64), 2, axis=1, halo=sortlast.required_halo(p, (64, 64, 64)))
    decoded_full = np.zeros_like(full)
    ops = []
    for r in range(2):
        o, e = stored[r]
        vol = vr.Volume(ctx, *e).set_cell_origin(o)
        vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))
        streams, origins = [], []
        for k in range(2):
            for i in range(2):
                brick = np.ascontiguousarray(full[k * B:(k + 1) * B, r * B:(r + 1) * B, i * B:(i + 1) * B])
                s, _ = vr.Stream.compress(brick, 1, 2)
                s.prepare().upload(ctx)
                streams.append(s)
                origins.append((i * B - o[0], r * B - o[1], k * B - o[2]))
                hd = s.header
                rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, B, B, B)
                assert rc == 0
                decoded_full[k * B:(k + 1) * B, r * B:(r + 1) * B, i * B:(i + 1) * B] = dec
        plan = vr.DecodePlan(ctx, streams, origins, vol)
        plan.run()
        ops.append(sortlast.CudaSlabOps(ctx, sortlast.SortLast(ctx, descs, r), vol, W, H))
        ops[-1]._keep = (streams, plan)
    sortlast.exchange_halos_lockstep(ops, descs, stored)
    torch.cuda.synchronize()
    for r in range(2):
        assert np.array_equal(ops[r].vol.download(), cut_slab(decoded_full, descs[r], stored[r]))
    frames = sortlast.render_frame_lockstep(ops, p)
    want, _, _ = oa.oracle_render(decoded_full, render.oracle_param_blob(p), 8)
    assert np.abs(frames[0].cpu().numpy() - want).max() < 2e-5
