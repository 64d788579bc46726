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
if mode in ("peer", "peer-tin"):
    def gather_handles(mine):
        allh = [None] * world
        dist.all_gather_object(allh, mine)
        return allh
    exch = sortlast.PeerExchange(ctx, rank, world, 1024, 1024, halo_bytes=sortlast.halo_staging_bytes(descs, stored), handles=gather_handles)
else:
    exch = sortlast.TorchDistExchange()

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e

if mode == "peer-tin":
    # the round-2 protocol (sortlast.PeerExchange.frame_stages_hybrid with t_in): T_seg stripes to their owners, T_in back,
    # pass 2 on T_in, terms to the stripe owners, sum + colour map with the finished stripes to rank 0 only
    from ctypes import byref, c_void_p
    from volumerenderer_b200._lib import check, lib
    names = ["decode", "halo", "partial", "stripes_put_wait", "tin_pass_wait", "resolve", "terms_put_wait", "finish_to0_wait"]
    acc = {n: [] for n in names}; host = []
    ex, P, n = exch, 1024 * 1024, world
    everyone = list(range(n))
    for it in range(25):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        e = [ev()]
        plan.run(); e.append(ev())
        ex.exchange_halos(ops, descs, stored); e.append(ev())
        cseg, tseg = ops.partial_composite(p); e.append(ev())
        s4 = 4 * P // n
        ex._put(ex.GATHER, [(j, ex.off_gather + rank * s4, tseg.data_ptr() + j * s4, s4) for j in everyone]); ex._wait(ex.GATHER, everyone); e.append(ev())
        off_tin = ex.off_red_in + 8 * P
        check(lib.vr_peer_tin_composite(ctx.h, ex.h, ex.TIN, ops.sl.h, byref(p), ex.off_gather, off_tin)); ex._wait(ex.TIN, everyone); e.append(ev())
        c = ops.resolve_composite_tin(p, ex._base() + off_tin, tseg, cseg); e.append(ev())
        sb = 8 * P // n
        ex._put(ex.RED_IN, [(j, ex.off_red_in + rank * sb, c.data_ptr() + j * sb, sb) for j in everyone]); ex._wait(ex.RED_IN, everyone); e.append(ev())
        check(lib.vr_peer_finish_composite_to(ctx.h, ex.h, ex.RED_OUT, byref(p), ex.off_red_in, ex.off_total, 0))
        if rank == 0:
            ex._wait(ex.RED_OUT, everyone)
        e.append(ev())
        host.append((time.perf_counter() - t0) * 1e3)
        torch.cuda.synchronize()
        if it >= 5:
            for k, nm in enumerate(names):
                acc[nm].append(e[k].elapsed_time(e[k + 1]))
    rng = range(0)
else:
    names = ["decode", "halo", "partial", "all_gather", "resolve", "all_reduce", "finish"]
    acc = {n: [] for n in names}; host = []
    rng = range(25)
for it in rng:
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
