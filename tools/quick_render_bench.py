"""Quick stage-2 measurement: 64-brick RM layer k=7 (2048x2048x128) decoded on the GPU, then
composite / isosurface frames at 1024^2, default and oblique cameras, CUDA-event timed."""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth

NB = int(sys.argv[1]) if len(sys.argv) > 1 else 64
W = H = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
ctx = vr.Context(0)
side = int(round(NB ** 0.5))


def make(b):
    i, j = b % side, b // side
    s, _ = vr.Stream.compress(synth.rm_brick(i, j, 7), 1, 2)
    return s.prepare()


t0 = time.time()
with ThreadPoolExecutor(max_workers=os.cpu_count() or 8) as ex:
    streams = list(ex.map(make, range(side * side)))
print("compress+prepare s", time.time() - t0, file=sys.stderr)
for s in streams:
    s.upload(ctx)
vol = vr.Volume(ctx, 256 * side, 256 * side, 128)
plan = vr.DecodePlan(ctx, streams, [((b % side) * 256, (b // side) * 256, 0) for b in range(side * side)], vol)
plan.run(); ctx.synchronize()
out = torch.empty((H, W, 4), dtype=torch.float32, device="cuda:0")
cnt = torch.zeros(1, dtype=torch.int64, device="cuda:0")
torch.cuda.synchronize()
for mode, mname in ((render.COMPOSITE, "composite"), (render.ISOSURFACE, "isosurface")):
    for vname, p in (("default", render.default_params(W, H, mode)), ("oblique", render.oblique_params(W, H, mode))):
        for _ in range(2):
            render.render_device(ctx, vol, p, out.data_ptr())
        ctx.synchronize()
        cnt.zero_(); torch.cuda.synchronize()
        render.render_device(ctx, vol, p, out.data_ptr(), None, cnt.data_ptr()); ctx.synchronize()
        S = int(cnt.item())
        ts = []
        for _ in range(5):
            ctx.timer_begin(1); render.render_device(ctx, vol, p, out.data_ptr()); ctx.timer_end(1)
            ts.append(ctx.timer_ms(1))
        ms = float(np.median(ts))
        print(json.dumps({"mode": mname, "view": vname, "W": W, "ms": ms, "fps": 1000 / ms, "samples": S,
                          "gsamples_s": S / ms / 1e6, "mean_rgba": out.mean(dim=(0, 1)).tolist()}))
