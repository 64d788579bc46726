"""Stage-1 A/B (not the contract bench): the persistent warp-specialised decode kernel against the classic one-CTA-per-tile
kernel on the three workloads the judge's roofline uses -- the bench's rm_yslab slab (95 % constant fills), the config-3
layer (64 mixing-layer bricks, 5.3 bits/voxel) and noise bricks (incompressible, 14.8 bits/voxel).  CUDA events through the
context timers, inputs larger than L2.  One JSON line per (workload, kernel, part).

  python tools/decode_ab.py [--iters 20] [--workloads slab,config3,noise] [--check] > gpurun_out/decode_ab.jsonl
"""
import argparse
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

import bench  # noqa: E402
import volumerenderer_b200 as vr  # noqa: E402
from volumerenderer_b200 import synth  # noqa: E402


def timed(ctx, fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    ctx.synchronize()
    ms = []
    for _ in range(iters):
        ctx.timer_begin(0); fn(); ctx.timer_end(0)
        ms.append(ctx.timer_ms(0))
    return float(np.median(ms)), float(np.min(ms))


def workload(ctx, name, threads):
    if name == "slab":
        bricks, streams, _ = bench.build_streams(vr, 0, threads)
        origins = [(i * 256, 0, k * 128) for (i, j, k) in bricks]
        dims = (2048, 256, 1920)
    elif name == "config3":
        ijk = [(i, j, 7) for j in range(8) for i in range(8)]
        with ThreadPoolExecutor(max_workers=threads) as ex:
            vols = list(ex.map(lambda a: synth.rm_brick(*a), ijk))
        streams = [vr.Stream.compress_device(ctx, v, 1, 2)[0] for v in vols]
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(lambda s: s.prepare(), streams))
        origins = [(i * 256, j * 256, 0) for (i, j, k) in ijk]
        dims = (2048, 2048, 128)
    else:
        n = 16
        with ThreadPoolExecutor(max_workers=threads) as ex:
            vols = list(ex.map(lambda b: synth.noise_brick(seed=0xB200 + b), range(n)))
        streams = [vr.Stream.compress_device(ctx, v, 1, 2)[0] for v in vols]
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(lambda s: s.prepare(), streams))
        origins = [((b % 4) * 256, (b // 4) * 256, 0) for b in range(n)]
        dims = (1024, 1024, 128)
    for s in {id(s): s for s in streams}.values():
        s.upload(ctx)
    return streams, origins, dims


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--workloads", default="slab,config3,noise")
    ap.add_argument("--check", action="store_true", help="compare the two kernels' slabs byte for byte")
    ap.add_argument("--fill-mode", type=int, default=2, help="persistent kernel: 0 tensor stores, 1 streaming stores of the fill warp, 2 tensor stores + evict-first")
    ap.add_argument("--kernels", default="persistent,classic")
    ap.add_argument("--tma-rows", type=int, default=8, help="rows per tensor-store box of the constant groups: 8 or 16 (anything else: the classic kernel)")
    ap.add_argument("--unit-store", type=int, default=0, help="persistent kernel: 0 choose, 1 tensor store per unit, 2 16-byte stores of the lanes")
    ap.add_argument("--parse-order", type=int, default=0, help="0 address order, 1 longest segments first")
    ap.add_argument("--split", type=int, default=0, help="persistent kernel: 1 = constant groups as a kernel of their own on a side stream")
    ap.add_argument("--probes", action="store_true", help="persistent kernel: two more timing probes (results are not valid voxels)")
    args = ap.parse_args()
    peak, _ = bench.peaks()
    vr.debug_set("fill_mode", args.fill_mode)
    vr.debug_set("tma_rows", args.tma_rows)
    vr.debug_set("parse_order", args.parse_order)
    vr.debug_set("unit_store", args.unit_store)
    vr.debug_set("decode_split", args.split)
    ctx = vr.Context(0)
    threads = min(32, os.cpu_count() or 8)
    for name in args.workloads.split(","):
        streams, origins, dims = workload(ctx, name, threads)
        vol = vr.Volume(ctx, *dims)
        ref = None
        for kernel in args.kernels.split(","):
            vr.debug_set("decode_kernel", 0 if kernel == "persistent" else 1)
            plan = vr.DecodePlan(ctx, streams, origins, vol)
            by = plan.bytes()
            nodes = sum(s.header.numActiveNodes for s in streams)
            for part, only in (("all", 0), ("parse_only", 1), ("fill_only", 2)) + ((("all_no_voxel_stores", 3), ("all_loads_from_l2", 4)) if args.probes and kernel == "persistent" else ()):
                vr.debug_set("decode_only", only)
                if kernel == "classic" and only:
                    plan.close()
                    plan = vr.DecodePlan(ctx, streams, origins, vol)  # the classic kernel filters its work list at creation
                ms, mn = timed(ctx, plan.run, args.iters)
                algo = by["stream"] + by["voxels"]
                print(json.dumps({"workload": name, "kernel": plan.kernel(), "fill_mode": args.fill_mode, "tma_rows": args.tma_rows, "parse_order": args.parse_order, "unit_store": args.unit_store, "split": args.split, "part": part, "ms": ms, "min_ms": mn,
                                  "gvoxel_s": by["voxels"] / ms / 1e6 if only == 0 else None,
                                  "algo_GBps": algo / ms / 1e6 if only == 0 else None,
                                  "frac_of_hbm_peak": algo / ms / 1e6 / peak if only == 0 else None,
                                  "bits_per_voxel": 2.0 * nodes / by["voxels"], "stream_bytes": by["stream"], "voxel_bytes": by["voxels"],
                                  "index_bytes": by["index"]}), flush=True)
            vr.debug_set("decode_only", 0)
            if args.check:
                plan.close()
                plan = vr.DecodePlan(ctx, streams, origins, vol)
                plan.run(); ctx.synchronize()
                got = vol.download()
                if ref is None:
                    ref = got
                else:
                    print(json.dumps({"workload": name, "check": "persistent == classic", "equal": bool(np.array_equal(ref, got))}), flush=True)
            plan.close()
        vr.debug_set("decode_kernel", 0)
        vol.close()
        for s in {id(s): s for s in streams}.values():
            s.close()


if __name__ == "__main__":
    main()
