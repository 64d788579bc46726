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
    dist.barrier()
    if rank == 0:
        print(f"SORTLAST_NCCL_OK world={world} worst_abs_err={worst:.3g}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
