#!/usr/bin/env python
"""bench.py -- contract benchmark of the hot path (kd-tree decode -> ray-march), one process per GPU.

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W  # the reference's CPU path on the host cores

Workload `rm_yslab` (BASELINE config 4 shape): the synthetic Richtmyer-Meshkov field on the
reference's brick grid (256x256x128 bricks, grid 8x8x15).  Rank r owns the y-slab j = r: 120 bricks =
2048x256x1920 voxels (0.94 GiB); N = 8 is exactly the full 2048x2048x1920 volume (weak scaling, no
data-path collective for decode).  A step = stage 1 (one batched decode launch of the rank's 120 streams
into its slab) + stage 2 (one 1024^2 colour-compositing frame).  At N = 1 the frame is the plain march of the
rank's slab; at N > 1 the texture is the N slabs side by side (2048 x 256N x 1920) and the frame is the
sort-last composite over NCCL (halo rows exchanged after the decode, segment partials, all-gather of
T_seg, resolve, all-reduce, colour map: volumerenderer_b200/sortlast.py).  Inputs are larger than L2
(the slab alone is 8x the 126 MB L2), so no explicit flush is needed between iterations.
metric: decoded GVoxel/s through the whole step (decode + frame); frames/s and G samples/s are extra keys.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BX, BY, BZ = 256, 256, 128
GI, GK = 8, 15
IMG = 1024


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.p.terminate()
        sel = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.05] or [r for (_, r) in self.rows[-3:]]
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in sel:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
                for nme, val in zip(names, f[2:6]):
                    if val.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples_in_region": len([1 for (t, _) in self.rows if t0 <= t <= t1 + 0.05])}


def slab_bricks(rank):
    """(i, j, k) of the rank's 120 bricks, k-major so constant layers interleave with the mixing layer."""
    return [(i, rank, k) for k in range(GK) for i in range(GI)]


def build_streams(vr, rank, threads, want_volumes=False, t=0):
    """Synthesise and compress the rank's bricks (CPU, untimed set-up).  Identical constant bricks are
    compressed once and share a stream (memoised deterministic set-up, not part of any timed region)."""
    from volumerenderer_b200 import synth
    bricks = slab_bricks(rank)
    const_cache, const_lock = {}, threading.Lock()

    def one(ijk):
        i, j, k = ijk
        cv = synth.rm_brick_constant_value(i, j, k, t)
        if cv is not None:
            with const_lock:
                if cv not in const_cache:
                    s, _ = vr.Stream.compress(synth.constant_brick(value=cv), 1, 2)
                    const_cache[cv] = s.prepare()
            return const_cache[cv], (synth.constant_brick(value=cv) if want_volumes else None)
        vol = synth.rm_brick(i, j, k, t)
        s, _ = vr.Stream.compress(vol, 1, 2)  # tolerance 1, 2 epochs: main.cpp:253-254
        return s.prepare(), (vol if want_volumes else None)

    with ThreadPoolExecutor(max_workers=threads) as ex:
        res = list(ex.map(one, bricks))
    return bricks, [r[0] for r in res], [r[1] for r in res]


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def frame_check_multi(args, ctx, vol, ops, descs, stored, frame, peers, p, rank, world, local):
    """N > 1, after the timed region: the sort-last frame the step draws (same kernels, same exchange, same decoded
    slabs) against the plain single-texture march -- the N = 1 kernel, itself parity-tested against the shader
    restatement in tests/ -- of the WHOLE texture, which rank 0 assembles from every rank's owned box (NCCL send /
    recv, outside every timed region).  Also reads every rank's peer-window error word (a wait that timed out).
    Returns the dict emitted as `frame_check`; raises SystemExit above 1/255 or after a time-out."""
    import torch
    import torch.distributed as dist
    import volumerenderer_b200 as vr
    from volumerenderer_b200 import render
    from volumerenderer_b200._lib import check, lib
    from ctypes import c_int64, c_void_p
    dev = torch.device("cuda", local)
    img = frame().clone()
    ctx.synchronize(); torch.cuda.synchronize()
    to = torch.tensor([max([int(bool(q.timed_out())) for q in peers if q is not None] or [0])], device=dev)
    dist.all_reduce(to, op=dist.ReduceOp.MAX)
    d, (so, se) = descs[rank], stored[rank]
    ext = [int(d.own_hi[a] - d.own_lo[a]) for a in range(3)]
    own = ops.pack_box([int(d.own_lo[a] - so[a]) for a in range(3)], ext)
    ctx.synchronize(); torch.cuda.synchronize()
    res = None
    if rank == 0:
        G = [int(v) for v in descs[0].global_]
        full = vr.Volume(ctx, *G)
        tmp = torch.empty_like(own)
        for r in range(world):
            src = own
            if r != 0:
                dist.recv(tmp, src=r)
                torch.cuda.synchronize()
                src = tmp
            lo = [int(descs[r].own_lo[a]) for a in range(3)]
            er = [int(descs[r].own_hi[a] - descs[r].own_lo[a]) for a in range(3)]
            check(lib.vr_volume_unpack_box(ctx.h, full.h, (c_int64 * 3)(*lo), (c_int64 * 3)(*er), c_void_p(src.data_ptr())))
            ctx.synchronize()
        ref = torch.empty((IMG, IMG, 4), dtype=torch.float32, device=dev)
        render.render_device(ctx, full, p, ref.data_ptr())
        ctx.synchronize(); torch.cuda.synchronize()
        diff = (img.view(IMG, IMG, 4) - ref).abs()
        diff = torch.where(torch.isnan(diff), torch.zeros_like(diff), diff)
        res = {"max_abs": float(diff.max().item()), "rows": IMG, "pixels_differing_above_1e-4": int((diff.amax(-1) > 1e-4).sum().item()),
               "timed_out": int(to.item()), "tolerance": 1.0 / 255.0,
               "against": f"plain single-texture march (march_kernel, N=1 path) of the whole {G[0]}x{G[1]}x{G[2]} texture gathered on rank 0",
               "nontrivial_pixels": int(((ref[..., :3] - 1.0).abs().amax(-1) > 1e-3).sum().item())}
        del full, tmp
    else:
        dist.send(own, dst=0)
    ok = torch.tensor([1 if (res is None or (res["max_abs"] <= 1.0 / 255.0 and res["timed_out"] == 0 and res["nontrivial_pixels"] > 0)) else 0],
                      device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) == 0 or int(to.item()) != 0:
        if rank == 0:
            sys.stderr.write(f"bench.py: frame_check FAILED: {res}\n")
        dist.barrier()
        raise SystemExit(3)
    return res


def _median_ms(ctx, fn, iters=7, warm=2, slot=15):
    for _ in range(warm):
        fn()
    ctx.synchronize()
    ms = []
    for _ in range(iters):
        ctx.timer_begin(slot); fn(); ctx.timer_end(slot)
        ms.append(ctx.timer_ms(slot))
    return float(np.median(ms))


def bench_extras(ctx, vr, render, vol, plan, threads, peak, local):
    """N = 1, after the headline region: further frames of the rm_yslab slab (isosurface,
    lazy isosurface_compressed straight from the streams, oblique camera), fetched samples beside march steps, and the
    stream-dense decode workloads -- the config-3 layer (64 bricks of k = 7, 5.3 bits/voxel: BASELINE config 3) and 16
    incompressible noise bricks -- with the same algorithmic-bytes formula and the same CUDA-event timing as `roofline`."""
    import torch
    from volumerenderer_b200 import synth
    dev = f"cuda:{local}"
    out = torch.empty((IMG, IMG, 4), dtype=torch.float32, device=dev)
    hit = torch.empty((IMG, IMG), dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)

    def count(v, p, fetched):
        vr.debug_set("count_fetched", 1 if fetched else 0)
        try:
            cnt.zero_(); torch.cuda.synchronize()
            render.render_device(ctx, v, p, out.data_ptr(), None, cnt.data_ptr()); ctx.synchronize()
            return int(cnt.item())
        finally:
            vr.debug_set("count_fetched", 0)

    def frames(v, tag_params):
        res = {}
        for tag, p in tag_params:
            ms = _median_ms(ctx, lambda: render.render_device(ctx, v, p, out.data_ptr(), hit.data_ptr()))
            steps, fetched = count(v, p, False), count(v, p, True)
            res[tag] = {"ms": ms, "frames_per_s": 1e3 / ms, "samples_per_frame": steps, "fetched_samples": fetched,
                        "gsamples_per_s": steps / ms / 1e6}
        return res

    plan.run(); ctx.synchronize()
    views = [("composite_reference", render.default_params(IMG, IMG, render.COMPOSITE)),
             ("isosurface_reference", render.default_params(IMG, IMG, render.ISOSURFACE)),
             ("composite_oblique", render.oblique_params(IMG, IMG, render.COMPOSITE)),
             ("isosurface_oblique", render.oblique_params(IMG, IMG, render.ISOSURFACE))]
    ex = {"frames_rm_yslab": frames(vol, views)}
    # isosurface_compressed: decode + frame in one call; lazy = only the tiles the frame can read are decoded
    pc = render.default_params(IMG, IMG, render.ISOSURFACE_COMPRESSED)
    st = {}
    full_ms = _median_ms(ctx, lambda: render.render_compressed_device(ctx, plan, pc, out.data_ptr(), hit.data_ptr()))
    lazy_ms = _median_ms(ctx, lambda: st.update(render.render_compressed_lazy_device(ctx, plan, pc, out.data_ptr(), hit.data_ptr())))
    plan.run(); ctx.synchronize()  # the slab is complete again
    ex["isosurface_compressed_rm_yslab"] = {"full_decode_plus_frame_ms": full_ms, "lazy_decode_plus_frame_ms": lazy_ms,
                                            "frames_per_s_lazy": 1e3 / lazy_ms, "tiles": st.get("tiles"), "tiles_written": st.get("tiles_written")}

    def dense(name, vols, origins, dims):
        t0 = time.time()
        streams = [vr.Stream.compress_device(ctx, v, 1, 2)[0] for v in vols]
        t_build = time.time() - t0
        t0 = time.time()
        with ThreadPoolExecutor(max_workers=threads) as pool:
            list(pool.map(lambda s: s.prepare(), streams))
        t_prep = time.time() - t0
        for s in streams:
            s.upload(ctx)
        v = vr.Volume(ctx, *dims)
        pl = vr.DecodePlan(ctx, streams, origins, v)
        by = pl.bytes()
        ms = _median_ms(ctx, pl.run, iters=11, warm=3)
        algo = by["stream"] + by["voxels"]
        nodes = sum(s.header.numActiveNodes for s in streams)
        r = {"workload": name, "kernel": "kd_decode6_kernel" if pl.kernel() == "persistent" else "kd_decode_tiles_kernel",
             "bound": "hbm", "achieved": algo / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": algo / ms / 1e6 / peak, "ms": ms,
             "algorithmic_bytes": algo, "stream_bytes": by["stream"], "voxel_bytes": by["voxels"], "index_bytes": by["index"],
             "bits_per_voxel": 2.0 * nodes / by["voxels"], "gvoxels_per_s": by["voxels"] / ms / 1e6, "bricks": len(streams),
             "setup": {"gpu_producer_s": t_build, "host_index_walk_s": t_prep, "threads": threads}}
        return r, streams, v, pl

    ijk = [(i, j, 7) for j in range(8) for i in range(8)]
    with ThreadPoolExecutor(max_workers=threads) as pool:
        vols = list(pool.map(lambda a: synth.rm_brick(*a), ijk))
    r3, streams3, v3, pl3 = dense("config-3 layer: 64 bricks 256x256x128 of k = 7 (2048x2048x128)", vols,
                                  [(i * BX, j * BY, 0) for (i, j, k) in ijk], (8 * BX, 8 * BY, BZ))
    del vols
    st3 = {}
    f_full = _median_ms(ctx, lambda: render.render_compressed_device(ctx, pl3, pc, out.data_ptr(), hit.data_ptr()))
    f_lazy = _median_ms(ctx, lambda: st3.update(render.render_compressed_lazy_device(ctx, pl3, pc, out.data_ptr(), hit.data_ptr())))
    pl3.run(); ctx.synchronize()
    f_march = _median_ms(ctx, lambda: render.render_device(ctx, v3, pc, out.data_ptr(), hit.data_ptr()))
    ex["config3"] = {"decode_ms": r3["ms"], "decode_gvoxels_per_s": r3["gvoxels_per_s"], "frame_ms": f_march,
                     "isosurface_compressed_full_ms": f_full, "isosurface_compressed_lazy_ms": f_lazy,
                     "frames_per_s_lazy": 1e3 / f_lazy, "tiles": st3.get("tiles"), "tiles_written": st3.get("tiles_written"),
                     "image": [IMG, IMG], "isoValue": "40/255"}
    # host index walk of one mixing-layer stream on one thread (what vr_stream_prepare costs for a legacy container)
    one = vr.Stream.from_memory(streams3[27].header, streams3[27].distance_map, streams3[27].bits)
    t0 = time.time(); one.prepare(); host_ms = (time.time() - t0) * 1e3
    # the same index emitted by the GPU producer from its own symbol offsets (csrc/kd_build.cu): producer time with and without it
    brick337 = synth.rm_brick(3, 3, 7)

    def producer_ms(with_index):
        vr.debug_set("producer_index", 1 if with_index else 0)
        try:
            ts = []
            for _ in range(3):
                t0 = time.time(); sx, _ = vr.Stream.compress_device(ctx, brick337, 1, 2); ts.append((time.time() - t0) * 1e3); sx.close()
            return float(np.median(ts))
        finally:
            vr.debug_set("producer_index", 1)
    p_with, p_without = producer_ms(True), producer_ms(False)
    ex["index_build"] = {"host_ms_per_brick": host_ms, "host_threads_per_brick": 1, "bricks_in_parallel": threads,
                         "brick": "rm brick (3, 3, 7)", "nodes": int(one.header.numActiveNodes),
                         "device_ms_per_brick": max(p_with - p_without, 0.0),
                         "device": "index emitted by the GPU stream producer from its own symbol offsets (no walk of the stream); "
                                   "producer alone %.2f ms, with index %.2f ms per brick; streams from elsewhere: host walk, or none at all "
                                   "for indexed containers (e2e.cold)" % (p_without, p_with)}
    one.close()
    pl3.close(); v3.close()
    for s in streams3:
        s.close()
    with ThreadPoolExecutor(max_workers=threads) as pool:
        vols = list(pool.map(lambda b: synth.noise_brick(seed=0xB200 + b), range(16)))
    rn, streamsn, vn, pln = dense("noise: 16 incompressible bricks 256x256x128 (1024x1024x128)", vols,
                                  [((b % 4) * BX, (b // 4) * BY, 0) for b in range(16)], (4 * BX, 4 * BY, BZ))
    pln.close(); vn.close()
    for s in streamsn:
        s.close()
    ex["roofline_dense"] = {"config3": r3, "noise": rn, "frac": min(r3["frac"], rn["frac"]),
                            "note": "same formula and events as `roofline` (stream read once + voxels written once, over the "
                                    "measured HBM copy peak); the parser is bound by instruction issue, not by HBM"}
    return ex


def run_ours(args):
    import torch
    import torch.distributed as dist
    import volumerenderer_b200 as vr
    from volumerenderer_b200 import render
    rank, world, local = dist_setup(args.gpus)
    for kv in args.debug:
        k, _, v = kv.partition("=")
        vr.debug_set(k, int(v))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    threads = max(1, (os.cpu_count() or 8) // max(1, world))
    # one non-default stream for everything (capturable into a CUDA graph); the library launches on torch's
    # current stream so NCCL collectives and kernels are ordered
    main_stream = torch.cuda.Stream()
    torch.cuda.set_stream(main_stream)
    ctx = vr.Context(local, cuda_stream=main_stream.cuda_stream)
    t_setup = time.time()
    bricks, streams, _ = build_streams(vr, rank, threads)
    uniq = {id(s): s for s in streams}
    # the timestep's stream images in ONE pinned buffer + ONE device buffer: a timestep is one H2D copy (vr_stream_arena)
    arena0 = vr.StreamArena(ctx, list(uniq.values()))
    arena0.upload(ctx)
    p = render.default_params(IMG, IMG, render.COMPOSITE)
    peer = None
    if world > 1:
        from volumerenderer_b200 import sortlast
        if args.tin_min_ranks is not None:
            sortlast.PeerExchange.T_IN_MIN_RANKS = args.tin_min_ranks
        G = (GI * BX, BY * world, GK * BZ)
        descs, stored = sortlast.slab_layout(G, world, axis=1, halo=sortlast.required_halo(p, G))
        so, se = stored[rank]
        vol = vr.Volume(ctx, *se).set_cell_origin(so)  # owned rows + one halo row per interior side
        origins = [(i * BX - so[0], rank * BY - so[1], k * BZ - so[2]) for (i, j, k) in bricks]
        sl = sortlast.SortLast(ctx, descs, rank)
        ops = sortlast.CudaSlabOps(ctx, sl, vol, IMG, IMG, count_samples=True)
        # exchange steps over peer memory (NVLink stores + flags, sortlast.PeerExchange); --exchange nccl = torch.distributed
        if args.exchange in ("peer", "peer-fused"):
            def gather_handles(mine):
                allh = [None] * world
                dist.all_gather_object(allh, mine)
                return allh
            peer = sortlast.PeerExchange(ctx, rank, world, IMG, IMG, halo_bytes=sortlast.halo_staging_bytes(descs, stored),
                                         handles=gather_handles, halo_buffers=2)
            exch = peer
        else:
            exch = sortlast.TorchDistExchange()
    else:
        vol = vr.Volume(ctx, GI * BX, BY, GK * BZ)
        origins = [(i * BX, 0, k * BZ) for (i, j, k) in bricks]
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    by = plan.bytes()
    out = torch.empty((IMG, IMG, 4), dtype=torch.float32, device=f"cuda:{local}")
    cnt = torch.zeros(1, dtype=torch.int64, device=f"cuda:{local}")
    host_img = torch.empty((IMG, IMG, 4), dtype=torch.float32).pin_memory()
    torch.cuda.synchronize()
    t_setup = time.time() - t_setup

    def frame():
        if world > 1:
            if peer is not None:
                peer.exchange_halos(ops, descs, stored)
                return peer.render_frame(ops, p, fused=args.exchange == "peer-fused")
            sortlast.exchange_halos(ops, descs, stored, rank)
            return sortlast.render_frame(ops, exch, p)
        render.render_device(ctx, vol, p, out.data_ptr())
        return out

    def step(slot=None):
        # slot != None: bracket each stage with its own CUDA-event pair (read back after the region)
        if slot is None:
            plan.run()
            frame()
        else:
            ctx.timer_begin(slot); plan.run(); ctx.timer_end(slot)
            ctx.timer_begin(slot + 1); frame(); ctx.timer_end(slot + 1)

    def barrier():
        ctx.synchronize(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ctx.synchronize(); torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step(1)
    # nvidia-smi needs a few hundred ms to start printing: launch it now so that it is sampling when the timed region runs
    sampler = ClockSampler(local) if rank == 0 else None
    # samples per frame (effective march steps reaching the fetch), counted once outside the timed region
    torch.cuda.synchronize()
    if world > 1:
        ops.samples.zero_(); frame(); torch.cuda.synchronize()
        t = ops.samples.clone(); dist.all_reduce(t)
        samples = int(t.item())  # fetches of all ranks, both passes
        ops.samples = None       # not counted inside the timed region
    else:
        cnt.zero_(); torch.cuda.synchronize()
        render.render_device(ctx, vol, p, out.data_ptr(), None, cnt.data_ptr()); ctx.synchronize()
        samples = int(cnt.item())

    # The step (decode launch, occupancy refresh, march / sort-last frame with its NCCL collectives) is
    # captured once into a CUDA graph and replayed: one host call per step instead of ~10-30.
    # (N > 1: only with the peer-memory exchange, whose steps are plain kernels with device-side epoch counters;
    # capturing the NCCL point-to-point halo exchange together with the collectives hung in this torch/NCCL
    # combination, so --exchange nccl launches eagerly.)
    graph = None
    if not args.no_graph and (world == 1 or peer is not None):
        try:
            barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=main_stream):
                step(None)
            graph = g
            graph.replay(); barrier()
        except Exception as e:  # stay measurable: eager launches
            graph = None
            sys.stderr.write(f"bench.py: CUDA graph capture failed ({type(e).__name__}: {e}); eager launches\n")
            torch.cuda.synchronize()
    # Pipelined steps (default): a time-varying volume is decoded into one of TWO slabs while the frame of the previous
    # step is drawn from the other, on a second, high-priority stream -- the decode's fills are HBM-bound, the march is
    # issue-bound and (N > 1) the frame spends much of its time waiting for peers, so the two overlap well.  Every step
    # still decodes its 120 bricks and draws one frame of a freshly decoded slab; a pair of steps is one graph replay.
    pair = None
    peerF = None
    inter = None
    schedule = "sequential"
    if args.pipeline and world > 1 and peer is not None and args.exchange == "peer" and args.schedule == "interleaved":
        # --schedule interleaved (measured slightly slower than two streams on 2 and 8 GPUs, profiles/r2_schedule_ab.md): ONE stream, the stages of frame(t) and decode(t+1) interleaved so that every device-side wait of the
        # exchange sits behind a long kernel of the other stage (sortlast.InterleavedSteps)
        try:
            vol2 = vr.Volume(ctx, *se).set_cell_origin(so)
            plan2 = vr.DecodePlan(ctx, streams, origins, vol2)
            ops2 = sortlast.CudaSlabOps(ctx, sl, vol2, IMG, IMG, count_samples=False)
            opsA = sortlast.CudaSlabOps(ctx, sl, vol, IMG, IMG, count_samples=False)
            inter = sortlast.InterleavedSteps(peer, [plan, plan2], [opsA, ops2], descs, stored, p, frame_to=0)
            barrier()
            inter.prologue()
            for _ in range(4):
                inter.step()
            barrier()
            if not args.no_graph:
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2, stream=main_stream):
                    inter.step(); inter.step()
                g2.replay(); barrier()
                pair = g2.replay
            else:
                pair = lambda: (inter.step(), inter.step())
            schedule = "interleaved"
        except Exception as e:
            pair, inter = None, None
            sys.stderr.write(f"bench.py: interleaved steps unavailable ({type(e).__name__}: {e}); sequential steps\n")
            torch.cuda.synchronize()
    elif args.pipeline and (world == 1 or peer is not None):
        schedule = "two-stream"
        try:
            f_stream = torch.cuda.Stream(priority=-1)
            ctxF = vr.Context(local, cuda_stream=f_stream.cuda_stream)
            vols = [vol, (vr.Volume(ctx, *se).set_cell_origin(so) if world > 1 else vr.Volume(ctx, GI * BX, BY, GK * BZ))]
            plans = [plan, vr.DecodePlan(ctx, streams, origins, vols[1])]
            outs = [out, torch.empty_like(out)]
            if world > 1:
                opsF = [sortlast.CudaSlabOps(ctxF, sortlast.SortLast(ctxF, descs, rank), v, IMG, IMG, count_samples=False) for v in vols]
                peerF = sortlast.PeerExchange(ctxF, rank, world, IMG, IMG, halo_bytes=sortlast.halo_staging_bytes(descs, stored),
                                              handles=gather_handles)

            def frame_of(k):
                with torch.cuda.stream(f_stream):
                    if world > 1:
                        peerF.exchange_halos(opsF[k], descs, stored)
                        return peerF.render_frame(opsF[k], p, fused=args.exchange == "peer-fused", frame_to=0)
                    render.render_device(ctxF, vols[k], p, outs[k].data_ptr())
                    return outs[k]

            def step_pair():
                ctxF.wait_for(ctx)                      # the frame stream joins (slab 1 was decoded by the previous pair)
                plans[0].run(); frame_of(1)             # decode slab 0  ||  frame of slab 1
                ctx.wait_for(ctxF); ctxF.wait_for(ctx)
                plans[1].run(); frame_of(0)             # decode slab 1  ||  frame of slab 0
                ctx.wait_for(ctxF)

            plans[1].run(); barrier()
            for _ in range(3):
                step_pair()
            barrier()
            if not args.no_graph:
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2, stream=main_stream):
                    step_pair()
                g2.replay(); barrier()
                pair = g2.replay
            else:
                pair = step_pair
        except Exception as e:
            pair = None
            schedule = "sequential"
            sys.stderr.write(f"bench.py: pipelined steps unavailable ({type(e).__name__}: {e}); sequential steps\n")
            torch.cuda.synchronize()
    barrier()
    t0 = time.time()
    ctx.timer_begin(0)
    todo = args.steps
    if pair is not None:
        for it in range(args.steps // 2):
            pair()
        todo = args.steps & 1
    if inter is not None and pair is not None:
        for it in range(todo):
            inter.step()
        todo = 0
        inter.drain()  # the last issued frame is waited for inside the timed region
    for it in range(todo):
        if graph is not None:
            graph.replay()
        else:
            step(None)
    ctx.timer_end(0)
    total_ms = ctx.timer_ms(0)
    barrier()
    t1 = time.time()
    # per-stage CUDA-event pairs on extra eager steps right after the timed region (events cannot be read
    # back from inside a captured graph); same kernels, same inputs
    n_probe = 7
    for k in range(n_probe):
        step(1 + 2 * k)
    dec_ms = [ctx.timer_ms(1 + 2 * k) for k in range(n_probe)]
    mar_ms = [ctx.timer_ms(2 + 2 * k) for k in range(n_probe)]
    barrier()
    clocks = sampler.stop(t0, t1) if sampler else None
    if world > 1:
        t = torch.tensor([total_ms], device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_step = total_ms / args.steps
    frame_check = frame_check_multi(args, ctx, vol, ops, descs, stored, frame, [peer, peerF], p, rank, world, local) if world > 1 else None
    voxels_rank = by["voxels"]
    voxels_all = voxels_rank * world

    # ---- end to end through the C ABI with host buffers, every step: H2D of the step's stream images
    # (pinned), decode, frame, D2H of the RGBA frame into pinned host memory.  Two legs:
    #  serial   : one timestep, upload -> decode -> frame -> read-back, one after the other;
    #  streaming: timesteps t=0 / t=1 alternate (BASELINE config 5 shape); a copy context uploads the next
    #             timestep's streams while the render context decodes and draws the current one
    #             (vr_context_wait_for orders the two CUDA streams; device images are double-buffered
    #             because each timestep owns its vr_stream objects).  This is the headline e2e.
    d2h = IMG * IMG * 16
    e2e_steps = max(4, min(args.steps, 20))
    e2e_steps += e2e_steps & 1  # both timesteps equally often

    def reduce_max(ms):
        if world > 1:
            tt = torch.tensor([ms], device=f"cuda:{local}")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())
        return ms

    def serial_step():
        arena0.upload(ctx)
        plan.run()
        img = frame()
        if rank == 0:
            ctx.download(host_img.data_ptr(), img.data_ptr(), d2h)
        ctx.synchronize()

    for _ in range(2):
        serial_step()
    barrier()
    te = time.time()
    for _ in range(e2e_steps):
        serial_step()
    torch.cuda.synchronize()
    serial_ms = reduce_max((time.time() - te) * 1e3 / e2e_steps)

    _, streams1, _ = build_streams(vr, rank, threads, t=1)
    uniq1 = {id(s): s for s in streams1}
    ctx_copy = vr.Context(local)  # its own CUDA stream
    arena1 = vr.StreamArena(ctx_copy, list(uniq1.values()))
    arena1.upload(ctx_copy)
    ctx_copy.synchronize()
    plans = [plan, vr.DecodePlan(ctx, streams1, origins, vol)]
    sets = [list(uniq.values()), list(uniq1.values())]
    arenas = [arena0, arena1]
    host_imgs = [host_img, torch.empty((IMG, IMG, 4), dtype=torch.float32).pin_memory()]
    h2d_sets = [a.nbytes for a in arenas]  # what one upload copies
    h2d = sum(h2d_sets) / 2.0

    def streaming(n):
        arenas[0].upload(ctx_copy)
        for i in range(n):
            cur, nxt = i & 1, 1 - (i & 1)
            ctx.wait_for(ctx_copy)      # this step's streams have arrived
            ctx_copy.wait_for(ctx)      # the decode that last read the other set's device images is done
            if i + 1 < n:
                arenas[nxt].upload(ctx_copy)   # ONE copy per timestep
            plans[cur].run()
            img = frame()
            if rank == 0:
                ctx.download(host_imgs[cur].data_ptr(), img.data_ptr(), d2h)
        ctx.synchronize(); ctx_copy.synchronize()

    streaming(4)
    barrier()
    te = time.time()
    streaming(e2e_steps)
    torch.cuda.synchronize()
    e2e_ms = reduce_max((time.time() - te) * 1e3 / e2e_steps)

    # ---- what the host side of this box delivers: every rank copies its timestep arena H2D at the same time, nothing else
    # running (the e2e legs above are bound by this link at every N)
    ctx_copy.synchronize(); ctx.synchronize()
    barrier()
    reps = 10
    tp = time.time()
    for _ in range(reps):
        arenas[0].upload(ctx_copy)
    ctx_copy.synchronize()
    my_gbs = arenas[0].nbytes * reps / (time.time() - tp) / 1e9
    if world > 1:
        tt = torch.tensor([my_gbs, -my_gbs, my_gbs], device=f"cuda:{local}", dtype=torch.float64)
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        h2d_probe = {"per_rank_gbs_max": float(mx[0].item()), "per_rank_gbs_min": float(-mx[1].item()), "aggregate_gbs": float(sm[2].item())}
    else:
        h2d_probe = {"per_rank_gbs_max": my_gbs, "per_rank_gbs_min": my_gbs, "aggregate_gbs": my_gbs}
    h2d_probe["bytes_per_copy"] = arenas[0].nbytes
    h2d_probe["note"] = "all ranks at once, 10 copies of the timestep arena (pinned) each, wall clock"

    # ---- cold end to end (N = 1): every step opens the timestep's containers from files (page cache), indexes them,
    # pins + uploads, plans, decodes, draws, reads the frame back -- nothing is reused between steps.  Two container
    # formats: the reference's .bin (vr_streams_load walks every stream on the host threads to build the decode index) and
    # the indexed container of csrc/kd_kdx.cpp (the index is in the file: no walk).
    cold = None
    if world == 1 and not args.no_extras:
        import shutil
        import tempfile
        tmp = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        try:
            files = {"bin": [], "kdx": []}
            maps = []
            for t, (ss, us) in enumerate(((streams, uniq), (streams1, uniq1))):
                order = list(us.keys())
                maps.append([order.index(id(x)) for x in ss])
                for fmt in files:
                    pl = []
                    for m, key in enumerate(order):
                        f = os.path.join(tmp, f"t{t}_{m}.{fmt}")
                        (us[key].save if fmt == "bin" else us[key].save_indexed)(f)
                        pl.append(f)
                    files[fmt].append(pl)

            def cold_step(fmt, t, tm):
                t0 = time.time()
                ss = vr.load_streams(files[fmt][t], ctx, threads)       # open (+ host walk for .bin), pin, H2D
                t1 = time.time()
                pl = vr.DecodePlan(ctx, [ss[m] for m in maps[t]], origins, vol)  # work lists (host) + uploads
                t2 = time.time()
                pl.run()
                img = frame()
                ctx.download(host_img.data_ptr(), img.data_ptr(), d2h)
                ctx.synchronize()
                t3 = time.time()
                pl.close()
                for x in ss:
                    x.close()
                tm.append((t1 - t0, t2 - t1, t3 - t2, time.time() - t0))

            cold = {}
            for fmt in ("kdx", "bin"):
                tm = []
                cold_step(fmt, 0, [])
                for i in range(4):
                    cold_step(fmt, i & 1, tm)
                a = np.median(np.asarray(tm), axis=0) * 1e3
                cold[{"kdx": "indexed_container", "bin": "reference_container"}[fmt]] = {
                    "ms_per_step": float(a[3]), "open_index_pin_upload_ms": float(a[0]), "plan_ms": float(a[1]), "decode_frame_d2h_ms": float(a[2]),
                    "value": voxels_all / float(a[3]) / 1e6, "unit": "GVoxel/s",
                    "file_bytes_per_step": int(np.mean([sum(os.path.getsize(f) for f in files[fmt][t]) for t in range(2)]))}
            cold["host_threads"] = threads
            cold["note"] = ("every step: vr_streams_load of the timestep's files (page cache) -> vr_decode_plan_create -> decode -> frame -> D2H; "
                            "nothing reused between steps; reference_container = VolumeKdtree::save files, index built by the host walk; "
                            "indexed_container = csrc/kd_kdx.cpp files, no walk")
        finally:
            shutil.rmtree(tmp, ignore_errors=True)

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    extras = None
    if world == 1 and not args.no_extras:
        extras = bench_extras(ctx, vr, render, vol, plan, threads, peak, local)
    d_ms, m_ms = float(np.median(dec_ms)), float(np.median(mar_ms))  # eager probe steps; the first ones absorb rank skew
    dec_bytes = by["stream"] + by["voxels"]
    mar_bytes = 8 * samples // world + 16 * IMG * IMG  # per rank (N > 1: an even share of all ranks' fetches)
    roof_dec = {"kernel": "kd_decode6_kernel" if plan.kernel() == "persistent" else "kd_decode_tiles_kernel", "bound": "hbm", "achieved": dec_bytes / d_ms / 1e6, "peak": peak, "unit": "GB/s",
                "frac": dec_bytes / d_ms / 1e6 / peak, "traffic": None, "ms": d_ms, "algorithmic_bytes": dec_bytes,
                "index_bytes": by["index"], "peak_source": peak_src,
                "write_fraction": by["voxels"] / dec_bytes,
                "note": "peak = measured read+write copy bandwidth; a write-only fill of full 128-byte lines reaches the same "
                        "figure on this part (tools/fill_probe.cu, profiles/r1_fill_probe.txt)"}
    roof_mar = {"kernel": "march_kernel<composite>" if world == 1 else "sort-last frame (halo + partial + all-gather + resolve + all-reduce + finish)", "bound": "hbm", "achieved": mar_bytes / m_ms / 1e6, "peak": peak, "unit": "GB/s",
                "frac": mar_bytes / m_ms / 1e6 / peak, "traffic": None, "ms": m_ms, "algorithmic_bytes": mar_bytes,
                "peak_source": peak_src}
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            t = json.load(open(tr))
            roof_dec["traffic"] = t.get("decode"); roof_mar["traffic"] = t.get("march")
            # not measured by this run: one `ncu --set full` capture per kernel of this build (tools/profile_step.sh), committed
            roof_dec["traffic_source"] = roof_mar["traffic_source"] = t.get("source")
            if t.get("march_ncu"):
                # what bounds the march is instruction issue and L1, not HBM (its DRAM traffic is a few % of the algorithmic
                # bytes): the HBM figure above is SURVEY 8d's uniformity number, these are the kernel's own limits
                roof_mar["bound_note"] = "issue / L1-bound kernel; the HBM fraction is SURVEY 8d's uniformity figure, not its roofline"
                roof_mar["ncu"] = t["march_ncu"]
            if t.get("decode_ncu"):
                roof_dec["ncu"] = t["decode_ncu"]
        except Exception:
            pass
    dominant = roof_dec if d_ms >= m_ms else roof_mar
    nodes = sum(s.header.numActiveNodes for s in streams)
    line = {
        "metric": "decoded_gvoxels_per_s", "value": voxels_all / ms_step / 1e6, "unit": "GVoxel/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8 decode / f32 march", "data": "synthetic",
        "config": {"workload": "rm_yslab: per GPU 120 bricks 256x256x128 (y-slab 2048x256x1920 of the synthetic "
                               "Richtmyer-Meshkov volume, tolerance 1, 2 epochs) decoded then one 1024x1024 "
                               "colour-compositing frame (reference camera); N=8 is the full 2048x2048x1920 volume",
                   "bricks_per_gpu": len(streams), "image": [IMG, IMG], "bits_per_voxel": 2.0 * nodes / voxels_rank,
                   "l2": "inputs+outputs per step (>= 1 GB) exceed the 126 MB L2; no explicit flush",
                   "multi_gpu": ("single slab, plain march" if world == 1 else
                                 f"texture = {world} y-slabs (2048x{256 * world}x1920); decode has no collective; frame = sort-last composite "
                                 + ("over peer-memory windows on NVLink (halo planes; T_seg stripes to their owners and T_in back, 2 x 4 MiB x "
                                    "(n-1)/n per rank; (C,A) terms to the stripe owners; finished stripes to rank 0"
                                    " in the timed steps)" if peer is not None else
                                    "over NCCL (halo send/recv, all-gather T_seg 4 MiB/rank, all-reduce (C,A) 8 MiB)")),
                   "step": {"two-stream": "pipelined: two slabs; step t decodes one while the frame of step t-1 is drawn from the other on a "
                                          "second stream; every step = 120 bricks decoded + 1 frame of a freshly decoded slab",
                            "interleaved": "interleaved on one stream: two slabs; step t draws the frame of the slab decoded by step t-1 and "
                                           "decodes the next one, stage by stage (pass 1 | decode | halo put | resolve | halo unpack | sum), so the "
                                           "exchange's waits sit behind the other stage's kernels; every step = 120 bricks decoded + halo "
                                           "exchange + 1 frame of a freshly decoded slab",
                            "sequential": "sequential: decode, then frame, on one stream"}[schedule if pair is not None else "sequential"],
                   "schedule": schedule if pair is not None else "sequential"},
        "stages": {"decode_ms": d_ms, "march_ms": m_ms, "decode_gvoxels_per_s": voxels_rank / d_ms / 1e6,
                   "frames_per_s": 1e3 / m_ms, "gsamples_per_s": samples / m_ms / 1e6, "samples_per_frame": samples},
        "roofline": dominant, "roofline_decode": roof_dec, "roofline_march": roof_mar,
        "e2e": {"value": voxels_all / e2e_ms / 1e6, "unit": "GVoxel/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                "mode": "streaming: timesteps t=0/t=1 alternate, H2D of the next timestep (one copy of its arena) overlaps decode+frame of the current",
                "h2d_gbs": h2d * world / e2e_ms / 1e6, "h2d_probe": h2d_probe,
                "serial": {"value": voxels_all / serial_ms / 1e6, "ms_per_step": serial_ms,
                           "h2d_bytes_per_step": int(h2d_sets[0])},
                "cold": cold},
        # own kernels per step: decode, 3 occupancy refresh, march; N > 1 over NCCL: partial, resolve, finish instead of march;
        # N > 1 over peer windows (rank 0, one neighbour): decode, halo put + wait + box reduce, 3 occupancy, partial, gather put +
        # wait, T_in pass + wait, resolve, terms put + wait, sum + colour map + stripes, wait = 19
        "debug_knobs": args.debug or None, "exchange": None if world == 1 else args.exchange, "cuda_graph": graph is not None, "pipelined": pair is not None, "frame_check": frame_check, "gpu_launches": (5 if world == 1 else (19 if peer is not None else 7)) * args.steps, "clocks": clocks, "setup_s": t_setup,
    }
    if extras:
        line["roofline_dense"] = extras.pop("roofline_dense")
        line["index_build"] = extras.pop("index_build")
        line["extras"] = extras
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(budget_s=20.0, steps=1)
    emit(line)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------
# CPU side: the reference's own levelCut (oracle/_ref, unmodified reference code) brick-parallel over
# the host cores + the scalar shader restatement (the reference has no CPU renderer) for the frame.
def _cpu_setup(rank, threads):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_api as oa
    import volumerenderer_b200 as vr  # compressor only (CPU, reference algorithm); never the decode path
    from volumerenderer_b200 import render
    bricks, streams, _ = build_streams(vr, rank, threads)
    handles = []  # one reference object per brick: levelCut keeps per-call state in members (not re-entrant)
    kind = "reference" if oa.REF is not None else "port"
    import tempfile
    paths = {}
    for s in streams:
        hd = s.header
        if kind == "reference":
            if id(s) not in paths:
                with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
                    paths[id(s)] = f.name
                s.save(paths[id(s)])  # container written by our save(), opened by the reference's open()
            handles.append(oa.REF.ref_kd_open(paths[id(s)].encode()))
        else:
            handles.append((s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth))
    for pth in paths.values():
        os.remove(pth)
    return oa, render, bricks, streams, handles, kind


def _cpu_decode(oa, kind, handles, streams, idx, threads, reps=1):
    """levelCut of bricks idx (one brick per thread at a time); returns seconds."""
    if kind == "reference":
        import ctypes
        arr = (ctypes.c_void_p * len(idx))(*[handles[b] for b in idx])
        return oa.REF.ref_kd_levelcut_bench(arr, len(idx), reps, threads)
    t0 = time.time()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(lambda b: oa.oracle_levelcut(*handles[b], BX, BY, BZ), idx * reps))
    return time.time() - t0


def _cpu_full_volume(oa, kind, handles, streams, bricks):
    """Host copy of the decoded slab (input of the CPU frame), decoded by the oracle, untimed."""
    slab = np.empty((GK * BZ, BY, GI * BX), np.uint8)
    cache = {}
    for b, (i, j, k) in enumerate(bricks):
        key = id(streams[b])
        if key not in cache:
            s = streams[b]
            hd = s.header
            rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, BX, BY, BZ)
            assert rc == 0
            cache[key] = dec
        slab[k * BZ:(k + 1) * BZ, :, i * BX:(i + 1) * BX] = cache[key]
    return slab


def cpu_baseline(budget_s=20.0, steps=1, warmup=0, as_line=False, args=None):
    threads = os.cpu_count() or 1
    oa, render, bricks, streams, handles, kind = _cpu_setup(0, threads)
    slab = _cpu_full_volume(oa, kind, handles, streams, bricks)
    p = render.default_params(IMG, IMG, render.COMPOSITE)
    blob = render.oracle_param_blob(p)
    img = np.zeros((IMG, IMG, 4), np.float32)
    import ctypes
    buf = ctypes.create_string_buffer(blob, len(blob))

    def one_step(m):
        idx = list(range(0, len(bricks), m))
        td = _cpu_decode(oa, kind, handles, streams, idx, threads)
        t0 = time.time()
        oa.ORACLE.oracle_render_rows(slab.ctypes.data, GI * BX, BY, GK * BZ, buf, img.ctypes.data, None, threads, m)
        tr = time.time() - t0
        return td, tr, len(idx)

    # choose the sample stride m so that (warmup + steps) passes fit the budget
    td, tr, _ = one_step(8)
    est_full = (td + tr) * 8
    m = 1
    while est_full / m * (steps + warmup) > budget_s and m < 8:
        m += 1
    for _ in range(warmup):
        one_step(m)
    tds, trs, nb = [], [], 0
    for _ in range(steps):
        td, tr, nb = one_step(m)
        tds.append(td); trs.append(tr)
    t_step = float(np.mean(tds) + np.mean(trs))
    vox = nb * BX * BY * BZ
    sample = (f"every {m}th brick of the rank-0 slab ({nb} of {len(bricks)} bricks, reference levelCut, one brick per thread) + "
              f"every {m}th row of the 1024x1024 composite frame (C++ shader restatement, rows over threads)")
    base = {"value": vox / t_step / 1e9, "unit": "GVoxel/s", "cores": threads, "kind": kind, "sample": sample,
            "decode_s": float(np.mean(tds)), "frame_s": float(np.mean(trs)),
            "decode_mvoxels_per_s": vox / float(np.mean(tds)) / 1e6, "frames_per_s_full_frame": 1.0 / (float(np.mean(trs)) * m)}
    if not as_line:
        return base
    line = {"impl": "reference", "metric": "decoded_gvoxels_per_s", "value": base["value"], "unit": "GVoxel/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8 decode / f32 march", "data": "synthetic",
            "config": {"workload": "rm_yslab (see the GPU arm): bounded sample of the rank-0 slab per step", "image": [IMG, IMG]},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "GVoxel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    return line


_RESULT_FD = None


def emit(obj):
    data = (json.dumps(obj) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "peer-fused", "nccl"],
                    help="N > 1: sort-last exchange over peer-memory windows (copy kernels, or fused into the march kernels) or NCCL")
    ap.add_argument("--no-pipeline", dest="pipeline", action="store_false",
                    help="strictly sequential steps (decode, then frame) instead of decode(t+1) overlapping frame(t) on two slabs")
    ap.add_argument("--schedule", default="two-stream", choices=["interleaved", "two-stream"],
                    help="N > 1 with the peer exchange: how decode(t+1) and frame(t) share the GPU (one stream, stages interleaved; or two streams)")
    ap.add_argument("--tin-min-ranks", type=int, default=None, help="A/B: rank count from which the T_in stripe exchange replaces the all-gather of T_seg")
    ap.add_argument("--debug", action="append", default=[], metavar="KEY=VALUE",
                    help="A/B runs: vr_debug_set(KEY, VALUE) before anything is created (e.g. decode_split=1, decode_only=2: timing probes)")
    ap.add_argument("--no-extras", action="store_true", help="N = 1: skip the extra frames and the stream-dense decode workloads")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of replaying a captured CUDA graph")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON result: libraries that print to fd 1 (NCCL's version banner, ...) are
    # sent to stderr for the duration of the run and the result is written to the saved descriptor at the end
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        emit(cpu_baseline(budget_s=150.0, steps=args.steps, warmup=args.warmup, as_line=True, args=args))
        return
    run_ours(args)


if __name__ == "__main__":
    main()
