"""Probe: pure-fill decode (constant bricks) at 1 GB vs a plain device memset, CUDA-event timed."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumerenderer_b200 as vr
from volumerenderer_b200 import synth
ctx = vr.Context(0)
n = 120
s, _ = vr.Stream.compress(synth.constant_brick(value=230), 1, 2); s.prepare(); s.upload(ctx)
for shape, origins in (((256, 256, 128 * n), [(0, 0, 128 * b) for b in range(n)]),
                       ((2048, 256, 1920), [((b % 8) * 256, 0, (b // 8) * 128) for b in range(n)])):
    vol = vr.Volume(ctx, *shape)
    plan = vr.DecodePlan(ctx, [s] * n, origins, vol)
    for _ in range(3): plan.run()
    ctx.synchronize()
    ts = []
    for _ in range(10):
        ctx.timer_begin(0); plan.run(); ctx.timer_end(0); ts.append(ctx.timer_ms(0))
    ms = float(np.median(ts)); by = plan.bytes()["voxels"]
    print(json.dumps({"what": "tile-fill decode", "slab": shape, "ms": ms, "GBps": by / ms / 1e6}))
    plan.close(); vol.close()
t = torch.empty(n * 8 * 2**20, dtype=torch.uint8, device="cuda")
for _ in range(3): t.fill_(7)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
ts = []
for _ in range(10):
    e0.record(); t.fill_(7); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms = float(np.median(ts))
print(json.dumps({"what": "torch fill_ 1 GB", "ms": ms, "GBps": t.numel() / ms / 1e6}))
