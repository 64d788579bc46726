"""Host-side logic of the sort-last renderer on CPU (no GPU): slab layout, halo plan, visibility orders
(C ABI, host-only), and the exchange protocol of volumerenderer_b200/sortlast.py run (a) lock-step in one
process and (b) over world_size-2 gloo, with the oracle's slab restatement standing in for the kernels.
Ground truth: the single-texture oracle render of the whole volume.  Tolerances: composite RGBA max-abs
<= 1/255 (the segment algebra reorders fp32 products, expected ~1e-6); isosurface identical."""
import os
import socket

import numpy as np
import pytest
import torch

import oracle_api as oa
from sortlast_cpu import OracleSlabOps, cut_slab
from volumerenderer_b200 import render, sortlast, synth

TOL = 1.0 / 255.0


def _volume():
    return synth.rm_region(640, 1000, 900, 64, 64, 96)  # (Z=96, Y=64, X=64): mixing layer + both fluids


def _views(W, H, mode):
    return [render.default_params(W, H, mode, step=(1 / 64, 1 / 64, 1 / 96)),
            render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 128)),
            render.default_params(W, H, mode, eye=(0.2, 0.9, -0.4), front=(-0.2, -0.9, 0.4), up=(0, 0, 1), step=(1 / 64, 1 / 64, 1 / 96))]


def test_layout_and_halo_plan():
    descs, stored = sortlast.slab_layout((2048, 2048, 1920), 8, axis=1, halo=(1, 1, 1))
    assert [tuple(d.own_lo) for d in descs][:2] == [(0, 0, 0), (0, 256, 0)]
    assert stored[0] == ([0, 0, 0], [2048, 257, 1920]) and stored[3] == ([0, 767, 0], [2048, 258, 1920])
    tr = sortlast.halo_transfers(descs, stored)
    assert len(tr) == 14 and all(abs(s - d) == 1 and e == [2048, 1, 1920] for s, d, lo, e in tr)
    p = render.default_params(64, 64, render.ISOSURFACE)
    assert sortlast.required_halo(p, (2048, 2048, 1920)) == [31, 31, 37]
    p.mode = render.COMPOSITE
    assert sortlast.required_halo(p, (2048, 2048, 1920)) == [1, 1, 1]


def test_visibility_orders():
    descs, _ = sortlast.slab_layout((64, 64, 32), 4, axis=1)
    sl = sortlast.SortLast(None, descs, 1)
    for octant in range(8):
        want = [3, 2, 1, 0] if octant & 2 else [0, 1, 2, 3]
        assert sl.order(octant) == want
    with pytest.raises(Exception):  # overlapping boxes
        d2, _ = sortlast.slab_layout((64, 64, 32), 2, axis=1)
        d2[1].own_lo[1] = 16
        sortlast.SortLast(None, d2, 0)


def _desc(glob, lo, hi):
    d = sortlast.SlabDesc()
    d.global_[:] = glob
    d.origin[:] = lo
    d.own_lo[:] = lo
    d.own_hi[:] = hi
    return d


def test_only_grid_partitions_are_accepted():
    """Disjoint boxes that tile the texture but are not a grid have no front-to-back order valid for every ray (a +x+y ray
    can leave Q = [1,2]x[0,1] and enter P = [0,2]x[1,2] although P sorts before Q): vr_sortlast_create must refuse them."""
    G = (64, 64, 32)
    bad = [_desc(G, (0, 32, 0), (64, 64, 32)), _desc(G, (32, 0, 0), (64, 32, 32)), _desc(G, (0, 0, 0), (32, 32, 32))]
    with pytest.raises(Exception) as e:
        sortlast.SortLast(None, bad, 0)
    assert "grid" in str(e.value)
    good = [_desc(G, (0, 0, 0), (32, 32, 32)), _desc(G, (32, 0, 0), (64, 32, 32)), _desc(G, (0, 32, 0), (32, 64, 32)),
            _desc(G, (32, 32, 0), (64, 64, 32))]
    sl = sortlast.SortLast(None, good, 0)
    assert sl.order(0) == [0, 2, 1, 3] or sl.order(0) == [0, 1, 2, 3]
    assert sl.order(3)[0] == 3 and sl.order(3)[-1] == 0


@pytest.mark.parametrize("n,axis", [(2, 1), (4, 1), (3, 2)])
@pytest.mark.parametrize("mode", [render.COMPOSITE, render.ISOSURFACE])
def test_lockstep_matches_single_texture_oracle(n, axis, mode):
    full = _volume()
    Z, Y, X = full.shape
    for p in _views(48, 36, mode):
        halo = sortlast.required_halo(p, (X, Y, Z))
        descs, stored = sortlast.slab_layout((X, Y, Z), n, axis=axis, halo=halo)
        ops = [OracleSlabOps(sortlast.SortLast(None, descs, r), descs[r], cut_slab(full, descs[r], stored[r], with_halo=False), 48, 36)
               for r in range(n)]
        sortlast.exchange_halos_lockstep(ops, descs, stored)
        for r in range(n):
            assert np.array_equal(ops[r].slab, cut_slab(full, descs[r], stored[r]))
        frames = sortlast.render_frame_lockstep(ops, p)
        want, _, _ = oa.oracle_render(full, render.oracle_param_blob(p), 4)
        for f in frames:
            got = f.numpy()
            if mode == render.ISOSURFACE:
                assert np.array_equal(got, want)
            else:
                assert np.abs(got - want).max() <= TOL
                assert np.abs(got - want).max() < 2e-5


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = _volume()
        Z, Y, X = full.shape
        res = []
        for mode in (render.COMPOSITE, render.ISOSURFACE):
            p = render.oblique_params(80, 60, mode, step=(1 / 96, 1 / 96, 1 / 128))
            halo = sortlast.required_halo(p, (X, Y, Z))
            descs, stored = sortlast.slab_layout((X, Y, Z), world, axis=1, halo=halo)
            ops = OracleSlabOps(sortlast.SortLast(None, descs, rank), descs[rank], cut_slab(full, descs[rank], stored[rank], with_halo=False), 80, 60, threads=2)
            sortlast.exchange_halos(ops, descs, stored, rank)
            assert np.array_equal(ops.slab, cut_slab(full, descs[rank], stored[rank]))
            frame = sortlast.render_frame(ops, sortlast.TorchDistExchange(), p)
            res.append(frame.numpy().copy())
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_matches_single_texture_oracle():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    out = dict(q.get(timeout=240) for _ in range(2))
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    full = _volume()
    for k, mode in enumerate((render.COMPOSITE, render.ISOSURFACE)):
        p = render.oblique_params(80, 60, mode, step=(1 / 96, 1 / 96, 1 / 128))
        want, _, _ = oa.oracle_render(full, render.oracle_param_blob(p), 4)
        for r in range(2):
            if mode == render.ISOSURFACE:
                assert np.array_equal(out[r][k], want)
            else:
                assert np.abs(out[r][k] - want).max() < 2e-5
