"""Per-stage timing of one sort-last step (torchrun, N >= 2): CUDA events around every piece of
bench.py's multi-GPU step, plus host time per step (launch-bound check).  Diagnostic only."""
import os, sys, time, json
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, sortlast

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = vr.Context(local, cuda_stream=torch.cuda.current_stream().cuda_stream)
bricks, streams, _ = bench.build_streams(vr, rank, max(1, (os.cpu_count() or 8) // world))
for s in {id(s): s for s in streams}.values():
    s.upload(ctx)
view = sys.argv[1] if len(sys.argv) > 1 else "default"
p = render.default_params(1024, 1024, render.COMPOSITE) if view == "default" else render.oblique_params(1024, 1024, render.COMPOSITE)
G = (2048, 256 * world, 1920)
descs, stored = sortlast.slab_layout(G, world, axis=1, halo=sortlast.required_halo(p, G))
so, se = stored[rank]
vol = vr.Volume(ctx, *se).set_cell_origin(so)
plan = vr.DecodePlan(ctx, streams, [(i * 256 - so[0], rank * 256 - so[1], k * 128 - so[2]) for (i, j, k) in bricks], vol)
ops = sortlast.CudaSlabOps(ctx, sortlast.SortLast(ctx, descs, rank), vol, 1024, 1024, count_samples=False)
mode = sys.argv[2] if len(sys.argv) > 2 else "nccl"
if mode == "peer":
    def gather_handles(mine):
        allh = [None] * world
        dist.all_gather_object(allh, mine)
        return allh
    exch = sortlast.PeerExchange(ctx, rank, world, 1024, 1024, halo_bytes=sortlast.halo_staging_bytes(descs, stored), handles=gather_handles)
else:
    exch = sortlast.TorchDistExchange()

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e

names = ["decode", "halo", "partial", "all_gather", "resolve", "all_reduce", "finish"]
acc = {n: [] for n in names}; host = []
for it in range(25):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    e = [ev()]
    plan.run(); e.append(ev())
    if mode == "peer":
        exch.exchange_halos(ops, descs, stored)
    else:
        sortlast.exchange_halos(ops, descs, stored, rank)
    e.append(ev())
    cseg, tseg = ops.partial_composite(p); e.append(ev())
    ta = exch.all_gather(tseg); e.append(ev())
    c = ops.resolve_composite(p, ta, cseg); e.append(ev())
    tot = exch.all_reduce_sum(c); e.append(ev())
    ops.finish_composite(p, tot); e.append(ev())
    host.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    if it >= 5:
        for k, n in enumerate(names):
            acc[n].append(e[k].elapsed_time(e[k + 1]))
res = {n: float(np.mean(v)) for n, v in acc.items()}
res["host_issue_ms"] = float(np.mean(host[5:])); res["sum_ms"] = sum(res[n] for n in names)
allr = [None] * world
dist.all_gather_object(allr, res)
if rank == 0:
    for r, x in enumerate(allr):
        print(f"rank {r} view {view} {mode}: " + " ".join(f"{k}={v:.3f}" for k, v in x.items()))
dist.barrier(); dist.destroy_process_group()
