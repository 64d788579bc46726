"""TEST SCRIPT (launched by torchrun on >= 2 GPUs, see test_gpu_sortlast_nccl.py): the sort-last frame over
real NCCL collectives and over peer-memory windows, one process per GPU, against the single-texture oracle of the whole volume."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_api as oa  # noqa: E402
import volumerenderer_b200 as vr  # noqa: E402
from volumerenderer_b200 import render, sortlast, synth  # noqa: E402


class _OwnedBoxWriter:
    """Stands in for a decode plan in interleaved_steps_check: run() rewrites the slab (owned box of `full`, halo zeroed)."""

    def __init__(self, vol, full, desc, origin, extent):
        self.vol, self.e = vol, extent
        lo, hi = list(desc.own_lo), list(desc.own_hi)
        self.box = np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]])
        self.at = [lo[a] - origin[a] for a in range(3)]
        self.zero = np.zeros((extent[2], extent[1], extent[0]), np.uint8)

    def run(self):
        self.vol.upload_brick(self.zero, (0, 0, 0))
        self.vol.upload_brick(self.box, self.at)


def interleaved_steps_check(ctx, rank, world, full):
    """sortlast.InterleavedSteps (frame(t) and decode(t+1) + halo(t+1) interleaved on one stream, two halo staging areas, the
    frame waited for one step later): two alternating volumes, every frame against the staged frame of the same volume."""
    Z, Y, X = full.shape
    W, H = 160, 120
    other = np.ascontiguousarray(full[::-1, :, ::-1])  # the second timestep
    p = render.oblique_params(W, H, render.COMPOSITE, step=(1 / 96, 1 / 96, 1 / 128))
    descs, stored = sortlast.slab_layout((X, Y, Z), world, axis=1, halo=sortlast.required_halo(p, (X, Y, Z)))
    o, e = stored[rank]

    def gather_handles(mine):
        allh = [None] * world
        dist.all_gather_object(allh, mine)
        return allh
    peer = sortlast.PeerExchange(ctx, rank, world, W, H, halo_bytes=sortlast.halo_staging_bytes(descs, stored), handles=gather_handles,
                                 halo_buffers=2)
    sl = sortlast.SortLast(ctx, descs, rank)
    vols = [vr.Volume(ctx, *e).set_cell_origin(o) for _ in range(2)]
    ops = [sortlast.CudaSlabOps(ctx, sl, v, W, H) for v in vols]
    plans = [_OwnedBoxWriter(vols[k], (full, other)[k], descs[rank], o, e) for k in range(2)]
    want = []
    for k in range(2):  # staged frames of the two timesteps
        plans[k].run()
        peer.exchange_halos(ops[k], descs, stored)
        want.append(peer.render_frame(ops[k], p).clone())
        torch.cuda.synchronize()
    dist.barrier()
    assert float((want[0] - want[1]).abs().max()) > 0.05, "the two timesteps must differ"
    inter = sortlast.InterleavedSteps(peer, plans, ops, descs, stored, p)
    got = []
    inter.prologue()
    for _ in range(7):
        inter.step(consume=lambda f: got.append(f.clone()))
    got.append(inter.drain().clone())
    torch.cuda.synchronize()
    assert not peer.timed_out(), hex(peer.last_error)
    assert len(got) == 7
    for t, f in enumerate(got):
        d = float((f - want[t & 1]).abs().max())
        assert d == 0.0, (rank, t, d)
    dist.barrier()
    # the finished frame delivered to rank 0 only (vr_peer_finish_composite_to): the other ranks neither receive nor wait
    inter = sortlast.InterleavedSteps(peer, plans, ops, descs, stored, p, frame_to=0)
    got = []
    inter.prologue()
    for _ in range(5):
        inter.step(consume=lambda f: got.append(f.clone()))
    last = inter.drain()
    torch.cuda.synchronize()
    assert not peer.timed_out(), hex(peer.last_error)
    if rank == 0:
        got.append(last.clone())
        assert len(got) == 5
        for t, f in enumerate(got):
            assert float((f - want[t & 1]).abs().max()) == 0.0, (rank, t)
    else:
        assert last is None and not got
    dist.barrier()
    # the same frames through the one-call C entries (csrc/vr_capi_frame.cu) on the same windows
    halo = sortlast.required_halo(p, (X, Y, Z))
    assert sortlast.SortLastFrame.window_bytes(descs, W, H, halo, 2) <= peer.bytes
    fr = sortlast.SortLastFrame(ctx, peer.h, sl, W, H, halo, halo_buffers=2)
    for k in range(2):
        plans[k].run()                       # wipes the halo
        fr.exchange_halos(vols[k], buffer=k)
        img = fr.render(vols[k], p).clone()
        torch.cuda.synchronize()
        assert float((img - want[k]).abs().max()) == 0.0, (rank, k)
    assert not peer.timed_out(), hex(peer.last_error)
    dist.barrier()
    fr.close()
    peer.close()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = vr.Context(local, cuda_stream=torch.cuda.current_stream().cuda_stream)
    full = synth.rm_region(640, 1000, 900, 64, 64, 96)
    Z, Y, X = full.shape
    W, H = 160, 120
    worst = 0.0
    for mode in (render.COMPOSITE, render.ISOSURFACE):
        for p in (render.default_params(W, H, mode, step=(1 / 64, 1 / 64, 1 / 96)),
                  render.oblique_params(W, H, mode, step=(1 / 96, 1 / 96, 1 / 128))):
            descs, stored = sortlast.slab_layout((X, Y, Z), world, axis=1, halo=sortlast.required_halo(p, (X, Y, Z)))
            o, e = stored[rank]
            vol = vr.Volume(ctx, *e).set_cell_origin(o)
            vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))
            lo, hi = list(descs[rank].own_lo), list(descs[rank].own_hi)
            vol.upload_brick(np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]), [lo[a] - o[a] for a in range(3)])
            ops = sortlast.CudaSlabOps(ctx, sortlast.SortLast(ctx, descs, rank), vol, W, H)
            sortlast.exchange_halos(ops, descs, stored, rank)
            torch.cuda.synchronize()
            assert np.array_equal(vol.download(), full[o[2]:o[2] + e[2], o[1]:o[1] + e[1], o[0]:o[0] + e[0]]), "halo exchange"
            got = sortlast.render_frame(ops, sortlast.TorchDistExchange(), p).cpu().numpy()
            # the same frame with the exchange over peer-memory windows (CUDA IPC + NVLink stores + flags)
            def gather_handles(mine):
                allh = [None] * world
                dist.all_gather_object(allh, mine)
                return allh
            peer = sortlast.PeerExchange(ctx, rank, world, W, H, halo_bytes=sortlast.halo_staging_bytes(descs, stored), handles=gather_handles)
            vol.upload_brick(np.zeros((e[2], e[1], e[0]), np.uint8), (0, 0, 0))  # wipe the halo, exchange it again over the windows
            vol.upload_brick(np.ascontiguousarray(full[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]), [lo[a] - o[a] for a in range(3)])
            peer.exchange_halos(ops, descs, stored)
            torch.cuda.synchronize()
            assert np.array_equal(vol.download(), full[o[2]:o[2] + e[2], o[1]:o[1] + e[1], o[0]:o[0] + e[0]]), "halo exchange over peer windows"
            for rep in range(4):
                # staged collectives over the windows, then (composite) the passes fused with their exchange step
                got_peer = peer.render_frame(ops, p, fused=rep >= 2)
                torch.cuda.synchronize()
                assert not peer.timed_out(), hex(peer.last_error)
                d = float(np.nanmax(np.abs(got_peer.cpu().numpy() - got)))
                assert d <= (2e-6 if mode == render.COMPOSITE else 0.0), (rank, mode, rep, d)
            dist.barrier()
            peer.close()
            want, _, _ = oa.oracle_render(full, render.oracle_param_blob(p), 4)
            err = float(np.nanmax(np.abs(got - want)))
            worst = max(worst, err)
            assert err <= (2e-5 if mode == render.COMPOSITE else 2.0 / 255.0), (rank, mode, err)
    interleaved_steps_check(ctx, rank, world, full)
    dist.barrier()
    if rank == 0:
        print(f"SORTLAST_NCCL_OK world={world} worst_abs_err={worst:.3g}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
