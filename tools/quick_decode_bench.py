"""Quick stage-1 measurement (not the contract bench): decode N bricks of one kind, CUDA-event timed."""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumerenderer_b200 as vr
from volumerenderer_b200 import synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ctx = vr.Context(0)


def make(kind, b):
    if kind == "rm":
        i, j, k = b % 8, (b // 8) % 8, 7
        vol = synth.rm_brick(i, j, k)
    elif kind == "noise":
        vol = synth.noise_brick(seed=0xB200 + b)
    else:
        vol = synth.constant_brick(value=b % 256)
    s, _ = vr.Stream.compress(vol, 1, 2)
    s.prepare()
    return s


for kind, n in (("rm", N), ("noise", max(2, N // 4)), ("constant", N)):
    t0 = time.time()
    with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 8)) as ex:
        streams = list(ex.map(lambda b: make(kind, b), range(n)))
    tprep = time.time() - t0
    for s in streams:
        s.upload(ctx)
    vol = vr.Volume(ctx, 256, 256, 128 * n)
    plan = vr.DecodePlan(ctx, streams, [(0, 0, 128 * b) for b in range(n)], vol)
    for _ in range(3):
        plan.run()
    ctx.synchronize()
    times = []
    for _ in range(10):
        ctx.timer_begin(0); plan.run(); ctx.timer_end(0)
        times.append(ctx.timer_ms(0))
    by = plan.bytes()
    ms = float(np.median(times))
    vox = by["voxels"]
    nodes = sum(s.header.numActiveNodes for s in streams)
    print(json.dumps({"kind": kind, "bricks": n, "ms": ms, "min_ms": min(times), "gvoxel_s": vox / ms / 1e6,
                      "algo_GBps": (by["stream"] + by["voxels"]) / ms / 1e6, "with_index_GBps": sum(by.values()) / ms / 1e6,
                      "bits_per_voxel": 2 * nodes / vox, "index_frac": by["index"] / (by["stream"] + by["voxels"]),
                      "host_compress_prepare_s": tprep}))
    plan.close(); vol.close()
    for s in streams:
        s.close()
