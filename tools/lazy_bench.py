"""Compressed-domain frames (BASELINE.json config 3 and the bench's rm_yslab slab): decode-everything-then-march
(vr_render_compressed) against the view-dependent lazy decode (vr_render_compressed_lazy).
One JSON line per (workload, mode, view); frames are checked bit-identical.  CUDA events via the context timers.

  python tools/lazy_bench.py [--iters 20] [--workloads slab,config3] > gpurun_out/lazy_bench.jsonl
"""
import argparse
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import volumerenderer_b200 as vr  # noqa: E402
from volumerenderer_b200 import render, synth  # noqa: E402


def timed(ctx, fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    ctx.synchronize()
    ms = []
    for _ in range(iters):
        ctx.timer_begin(0); fn(); ctx.timer_end(0)
        ms.append(ctx.timer_ms(0))
    return float(np.median(ms)), float(np.min(ms))


def run(ctx, name, streams, origins, dims, img, iters, emit):
    vol = vr.Volume(ctx, *dims)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    by = plan.bytes()
    out = torch.zeros((img, img, 4), dtype=torch.float32, device="cuda:0")
    out2 = torch.zeros_like(out)
    hit = torch.zeros((img, img), dtype=torch.int32, device="cuda:0")
    hit2 = torch.zeros_like(hit)
    torch.cuda.synchronize()
    for mode, mname in ((render.ISOSURFACE_COMPRESSED, "isosurface_compressed"), (render.COMPOSITE, "composite")):
        for vname, p in (("reference", render.default_params(img, img, mode)), ("oblique", render.oblique_params(img, img, mode))):
            stats = {}

            def lazy():
                stats.update(render.render_compressed_lazy_device(ctx, plan, p, out.data_ptr(), hit.data_ptr()))

            def full():
                render.render_compressed_device(ctx, plan, p, out2.data_ptr(), hit2.data_ptr())

            def decode_only():
                plan.run()

            full_ms, full_min = timed(ctx, full, iters)
            dec_ms, _ = timed(ctx, decode_only, iters)
            lazy_ms, lazy_min = timed(ctx, lazy, iters)
            full()
            ctx.synchronize()
            same = bool(torch.equal(out, out2) and torch.equal(hit, hit2))
            emit({"workload": name, "mode": mname, "view": vname, "image": [img, img], "bricks": len(streams),
                  "full_ms": full_ms, "full_min_ms": full_min, "full_decode_ms": dec_ms, "lazy_ms": lazy_ms, "lazy_min_ms": lazy_min,
                  "speedup": full_ms / lazy_ms, "tiles": stats["tiles"], "tiles_written": stats["tiles_written"],
                  "voxel_bytes_full": by["voxels"], "voxel_bytes_lazy": stats["voxel_bytes"], "stream_bytes_full": by["stream"],
                  "stream_bytes_lazy": stats["stream_bytes"], "frames_bit_identical": same,
                  "hits": int((hit2 >= 0).sum().item())})
    plan.close()
    vol.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--workloads", default="slab,config3")
    args = ap.parse_args()
    ctx = vr.Context(0)
    BX, BY, BZ = synth.BRICK
    threads = os.cpu_count() or 8

    def emit(o):
        print(json.dumps(o), flush=True)

    wl = args.workloads.split(",")
    if "slab" in wl:  # the bench's N = 1 workload: y-slab 2048 x 256 x 1920, 120 bricks
        bricks, streams, _ = bench.build_streams(vr, 0, threads)
        for s in {id(s): s for s in streams}.values():
            s.upload(ctx)
        origins = [(i * BX, 0, k * BZ) for (i, j, k) in bricks]
        run(ctx, "rm_yslab (120 bricks, 2048x256x1920)", streams, origins, (8 * BX, BY, 15 * BZ), 1024, args.iters, emit)
    if "config3" in wl:  # 64 bricks of layer k = 7 as an 8 x 8 x 1 grid
        ijk = [(i, j, 7) for j in range(8) for i in range(8)]
        with ThreadPoolExecutor(max_workers=threads) as ex:
            vols = list(ex.map(lambda a: synth.rm_brick(*a), ijk))
        streams = []
        for v in vols:
            s, _ = vr.Stream.compress_device(ctx, v, 1, 2)
            s.upload(ctx)
            streams.append(s)
        origins = [(i * BX, j * BY, 0) for (i, j, _) in ijk]
        run(ctx, "config3 (64 bricks of layer k=7, 2048x2048x128)", streams, origins, (8 * BX, 8 * BY, BZ), 1024, args.iters, emit)


if __name__ == "__main__":
    main()
