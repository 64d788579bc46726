"""Stream producer timings on the GPU box: GPU producer (csrc/kd_build.cu) vs the host compressor (kd_compress.cpp)
vs the UNMODIFIED reference build() (oracle/_ref, useThreads as main.cpp does), same bricks, streams compared.
Usage: python tests/producer_bench.py [reps]      (prints one JSON line per brick kind)
TEST INFRASTRUCTURE (lives under tests/ because it times the oracle's reference build next to the product's producers)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import volumerenderer_b200 as vr  # noqa: E402
from volumerenderer_b200 import synth  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    try:
        import oracle_api as oa  # test infrastructure: only used here as the timed CPU reference
    except Exception:
        oa = None
    ctx = vr.Context(0)
    kinds = {"rm": lambda: synth.rm_brick(3, 4, 7), "noise": synth.noise_brick, "constant": synth.constant_brick}
    for kind, mk in kinds.items():
        vol = np.ascontiguousarray(mk())
        Z, Y, X = vol.shape
        v = vr.Volume(ctx, X, Y, Z)
        v.upload_brick(vol, (0, 0, 0))
        ctx.synchronize()
        vr.Stream.compress_volume(ctx, v, (0, 0, 0), (X, Y, Z), 1, 2)  # warm-up: scratch allocation
        ts, iters = [], None
        for _ in range(reps):
            t0 = time.perf_counter()
            s, st, iters = vr.Stream.compress_volume(ctx, v, (0, 0, 0), (X, Y, Z), 1, 2)
            ts.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        h, _ = vr.Stream.compress(vol, 1, 2)
        t_host = time.perf_counter() - t0
        same = bool(np.array_equal(h.bits, s.bits) and bytes(h.distance_map) == bytes(s.distance_map))
        t_ref = None
        if oa is not None and getattr(oa, "REF", None) is not None:
            t0 = time.perf_counter()
            rs = oa.ref_build(vol, 1, 2, 1)
            t_ref = time.perf_counter() - t0
            same = same and bool(np.array_equal(np.asarray(rs["bits"])[: len(s.bits)], s.bits))
            oa.REF.ref_kd_free(rs["h"])
        nodes = int(s.header.numActiveNodes)
        print(json.dumps({"brick": kind, "voxels": X * Y * Z, "nodes": nodes, "bits_per_voxel": 2.0 * nodes / (X * Y * Z),
                          "gpu_producer_ms": 1e3 * float(np.median(ts)), "gpu_producer_min_ms": 1e3 * float(min(ts)),
                          "gpu_mvoxel_per_s": X * Y * Z / float(np.median(ts)) / 1e6,
                          "host_compressor_ms": 1e3 * t_host, "reference_build_ms": None if t_ref is None else 1e3 * t_ref,
                          "fixed_point_passes_total_worst": list(iters), "streams_identical": same}))


if __name__ == "__main__":
    main()
