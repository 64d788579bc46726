"""GPU test (-m gpu) of the peer-memory exchange (vr_peer_*, sortlast.PeerExchange): N virtual ranks on ONE device,
each with its own context and stream and driven stage by stage, windows opened by raw pointer (the CUDA-IPC path needs one process per rank and
is covered by tests/test_gpu_sortlast_nccl.py on >= 2 GPUs).  The frame through PeerExchange must equal the frame whose
collectives are done locally in lock-step: all-gather and min are exact, the sum is taken in rank order in both."""
import ctypes

import numpy as np
import pytest
import torch

import volumerenderer_b200 as vr
from sortlast_cpu import cut_slab
from volumerenderer_b200 import render, sortlast, synth

pytestmark = pytest.mark.gpu
W, H = 160, 128


@pytest.fixture(scope="module")
def full():
    return synth.rm_region(640, 1000, 900, 64, 64, 96)


def _ranks(full, n, axis, p):
    Z, Y, X = full.shape
    halo = sortlast.required_halo(p, (X, Y, Z))
    descs, stored = sortlast.slab_layout((X, Y, Z), n, axis=axis, halo=halo)
    ctxs, ops = [], []
    for r in range(n):
        c = vr.Context(0)  # own non-blocking stream: a rank's device-side wait must not hold up the others
        o, e = stored[r]
        vol = vr.Volume(c, *e).set_cell_origin(o)
        lo = [descs[r].own_lo[a] for a in range(3)]; hi = [descs[r].own_hi[a] for a in range(3)]
        vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))
        vol.upload_brick(np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]), [lo[a] - o[a] for a in range(3)])
        c.synchronize()
        ctxs.append(c)
        ops.append(sortlast.CudaSlabOps(c, sortlast.SortLast(c, descs, r), vol, W, H, count_samples=False))
    return descs, stored, ctxs, ops


def _peers(ctxs, n, halo_bytes):
    ex = [sortlast.PeerExchange(ctxs[r], r, n, W, H, halo_bytes=halo_bytes) for r in range(n)]
    wins = [e.raw() for e in ex]
    for e in ex:
        e.open_raw(wins)
    return ex


@pytest.mark.parametrize("n,axis", [(2, 1), (4, 1), (8, 1)])
@pytest.mark.parametrize("mode", [render.COMPOSITE, render.ISOSURFACE])
def test_frame_over_peer_windows(full, n, axis, mode):
    p = render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 128))
    if mode == render.ISOSURFACE:
        p.isoValue = 100 / 255
    descs, stored, ctxs, ops = _ranks(full, n, axis, p)
    ex = _peers(ctxs, n, halo_bytes=4 * 64 * 96 * 8)
    torch.cuda.synchronize()
    # halos through the windows: every rank issues its puts and its device-side wait without blocking the host
    pending = [ex[r].exchange_halos(ops[r], descs, stored, defer_receive=True) for r in range(n)]
    for r in range(n):
        ex[r].receive_halos(ops[r], pending[r])
    for c in ctxs:
        c.synchronize()
    assert not any(e.timed_out() for e in ex)
    for r in range(n):
        assert np.array_equal(ops[r].vol.download(), cut_slab(full, descs[r], stored[r])), "halo exchange over peer windows"
    # reference: same kernels, collectives done locally
    want = [f.clone() for f in sortlast.render_frame_lockstep(ops, p)]
    torch.cuda.synchronize()
    for rep in range(3):  # windows are reused frame after frame
        frames = sortlast.render_frame_peer_lockstep(ops, ex, p)
        for c in ctxs:
            c.synchronize()
        frames = [f.clone() for f in frames]
        assert not any(e.timed_out() for e in ex), [hex(e.last_error) for e in ex]
        for r in range(n):
            if mode == render.ISOSURFACE:  # one non-zero term per pixel: exact
                assert torch.equal(frames[r], want[r]), (rep, r)
            else:  # the ranks' terms are summed in rank order here, in torch's reduction order in the lock-step frame
                assert (frames[r] - want[r]).abs().max().item() <= 2e-6, (rep, r)
    if mode == render.COMPOSITE:
        # the exchange fused into the march kernels: same arithmetic as the staged peer frame, bit for bit
        staged = [f.clone() for f in frames]
        for rep in range(4):
            fused = sortlast.render_frame_fused_lockstep(ops, ex, p, hybrid=rep % 2 == 1)
            for c in ctxs:
                c.synchronize()
            assert not any(e.timed_out() for e in ex), [hex(e.last_error) for e in ex]
            for r in range(n):
                assert torch.equal(fused[r], staged[r]), (rep, r)
    for e in ex:
        e.close()
    for c in ctxs:
        c.close()


@pytest.mark.parametrize("n", [1, 2, 4, 8])
@pytest.mark.parametrize("mode", [render.COMPOSITE, render.ISOSURFACE])
def test_whole_frame_through_the_c_entry(full, n, mode):
    """vr_sortlast_exchange_halos + vr_sortlast_render (csrc/vr_capi_frame.cu): one C call per timestep / per frame, every
    rank on its own stream and NOT driven in lock-step (a rank's device-side wait is queued before the peers' puts are even
    launched) -- against the lock-step frame of the same kernels, and n = 1 against the plain march."""
    p = render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 128))
    if mode == render.ISOSURFACE:
        p.isoValue = 100 / 255
    Z, Y, X = full.shape
    halo = sortlast.required_halo(p, (X, Y, Z))
    descs, stored, ctxs, ops = _ranks(full, n, 1, p)
    want = None
    if n > 1:  # the lock-step frame needs complete halos; wipe them again afterwards so that the C entry has to fill them
        sortlast.exchange_halos_lockstep(ops, descs, stored)
        want = [f.clone() for f in sortlast.render_frame_lockstep(ops, p)]
        torch.cuda.synchronize()
        for r in range(n):
            o, e = stored[r]
            lo = [descs[r].own_lo[a] for a in range(3)]; hi = [descs[r].own_hi[a] for a in range(3)]
            ops[r].vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))
            ops[r].vol.upload_brick(np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]), [lo[a] - o[a] for a in range(3)])
            ctxs[r].synchronize()
    nbytes = sortlast.SortLastFrame.window_bytes(descs, W, H, halo, halo_buffers=2)
    peers = []
    for r in range(n):
        h = vr._lib.c_void_p()
        vr._lib.check(vr._lib.lib.vr_peer_create(ctxs[r].h, r, n, nbytes, ctypes.byref(h)))
        peers.append(h)
    wins = []
    for h in peers:
        w = vr._lib.c_void_p()
        vr._lib.check(vr._lib.lib.vr_peer_raw(h, ctypes.byref(w)))
        wins.append(w.value)
    for h in peers:
        vr._lib.check(vr._lib.lib.vr_peer_open_raw(h, (vr._lib.c_void_p * n)(*wins)))
    frames = [sortlast.SortLastFrame(ctxs[r], peers[r], ops[r].sl, W, H, halo, halo_buffers=2) for r in range(n)]
    for rep in range(3):
        for r in range(n):
            frames[r].exchange_halos(ops[r].vol, buffer=rep & 1)
        got = [frames[r].render(ops[r].vol, p) for r in range(n)]
        for c in ctxs:
            c.synchronize()
        for r in range(n):
            err = vr._lib.c_int32(0)
            vr._lib.check(vr._lib.lib.vr_peer_error(peers[r], ctypes.byref(err)))
            assert err.value == 0, hex(err.value)
            assert np.array_equal(ops[r].vol.download(), cut_slab(full, descs[r], stored[r])), "halo exchange through the C entry"
            if n == 1:
                plain = render.render_to_host(ctxs[0], ops[0].vol, p)[0]
                assert np.abs(got[0].cpu().numpy().reshape(plain.shape) - plain).max() <= (2e-6 if mode == render.COMPOSITE else 0.0)
            elif mode == render.ISOSURFACE:
                assert torch.equal(got[r], want[r]), (rep, r)
            else:
                assert (got[r] - want[r]).abs().max().item() <= 2e-6, (rep, r)
    for f in frames:
        f.close()
    for h in peers:
        vr._lib.lib.vr_peer_destroy(h)
    for c in ctxs:
        c.close()


def test_wait_without_peer_times_out_instead_of_hanging():
    c = vr.Context(0)
    e = sortlast.PeerExchange(c, 0, 2, 64, 64)
    e.open_raw([e.raw(), e.raw()])
    e._wait(e.GATHER, [1])  # nobody raises that flag
    assert e.timed_out()
    e.close(); c.close()
