"""GPU parity (-m gpu): the GPU stream producer (csrc/kd_build.cu) against the host compressor
(kd_compress.cpp, pinned to the unmodified reference build() in test_host_codec.py) -- byte-identical streams,
identical distance maps and statistics -- and against the reference's golden vectors."""
import glob
import os
import time

import numpy as np
import pytest

import volumerenderer_b200 as vr
from volumerenderer_b200 import synth

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


def _same(a, sa, b, sb):
    ha, hb = a.header, b.header
    assert ha.numActiveNodes == hb.numActiveNodes
    assert bytes(a.distance_map) == bytes(b.distance_map)
    assert np.array_equal(a.bits, b.bits)
    for f in ("maxErrorBeforeGrowth", "maxError", "meanL1", "meanL2", "zeroRunRewrites"):
        assert getattr(sa, f) == getattr(sb, f), f


def _small(kind, X, Y, Z):
    z, y, x = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    if kind == "noise":
        return np.random.RandomState(11).randint(0, 256, (Z, Y, X)).astype(np.uint8)
    if kind == "constant":
        return np.full((Z, Y, X), 77, np.uint8)
    if kind == "ramp":
        return ((x * 3 + y * 2 + z * 5) % 256).astype(np.uint8)
    h = Z / 2 + (Z / 6) * np.sin(2 * np.pi * x / X) * np.cos(2 * np.pi * y / Y)
    d = (z - h) / max(Z / 10, 1)
    rng = np.random.RandomState(7)
    v = np.where(d <= -1, 0, np.where(d >= 1, 230, np.clip(np.round(115 * (1 + d)) + rng.randint(-16, 17, d.shape) * (1 - np.abs(d)), 0, 255)))
    return v.astype(np.uint8)


@pytest.mark.parametrize("kind", ["rm", "noise", "ramp", "constant"])
@pytest.mark.parametrize("shape", [(64, 64, 32), (16, 8, 4), (128, 64, 64)])
@pytest.mark.parametrize("tol,epochs", [(1, 2), (6, 1), (0, 5)])
def test_gpu_producer_matches_host_small(ctx, kind, shape, tol, epochs):
    X, Y, Z = shape
    vol = _small(kind, X, Y, Z)
    a, sa = vr.Stream.compress_device(ctx, vol, tol, epochs)
    b, sb = vr.Stream.compress(vol, tol, epochs)
    _same(a, sa, b, sb)


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_gpu_producer_reproduces_reference_golden(ctx, path, tmp_path):
    """tests/golden/*.npz: input volume + the container written by the UNMODIFIED reference's build() + save()
    (oracle/make_golden.py).  The GPU producer's saved file must be the same bytes."""
    g = np.load(path)
    vol = g["volume"]
    a, _ = vr.Stream.compress_device(ctx, vol, int(g["tolerance"]), int(g["maxEpochs"]))
    p = tmp_path / "gpu.bin"
    a.save(p)
    assert p.read_bytes() == g["container"].tobytes()


@pytest.mark.parametrize("kind", ["rm", "noise", "constant", "rm_empty"])
def test_gpu_producer_full_size_brick(ctx, kind):
    """BASELINE brick shape 256x256x128; also: the stream decodes back within the tolerance."""
    vol = {"rm": lambda: synth.rm_brick(3, 4, 7), "noise": synth.noise_brick, "constant": synth.constant_brick,
           "rm_empty": lambda: synth.rm_brick(0, 0, 0)}[kind]()
    t0 = time.time()
    a, sa = vr.Stream.compress_device(ctx, vol, 1, 2)
    t1 = time.time()
    b, sb = vr.Stream.compress(vol, 1, 2)
    t2 = time.time()
    print(f"\n{kind}: GPU producer {t1 - t0:.3f} s, host compressor {t2 - t1:.3f} s, {a.header.numActiveNodes} nodes")
    _same(a, sa, b, sb)


def test_compress_volume_box_roundtrip(ctx):
    """A brick that lives inside a device volume: decode two bricks into a slab, re-compress one of them in place
    (no host round trip of voxels), and get the identical stream back."""
    bricks = [_small("rm", 64, 64, 32), _small("noise", 64, 64, 32)]
    streams = [vr.Stream.compress(b, 1, 2)[0] for b in bricks]
    vol = vr.Volume(ctx, 128, 64, 32)
    for s in streams:
        s.upload(ctx)
    plan = vr.DecodePlan(ctx, streams, [(0, 0, 0), (64, 0, 0)], vol)
    plan.run(); ctx.synchronize()
    dec = vol.download()
    for k, s in enumerate(streams):
        box = np.ascontiguousarray(dec[:, :, 64 * k:64 * k + 64])
        again, _, iters = vr.Stream.compress_volume(ctx, vol, (64 * k, 0, 0), (64, 64, 32), 1, 2)
        host, _ = vr.Stream.compress(box, 1, 2)
        assert again.header.numActiveNodes == host.header.numActiveNodes
        assert np.array_equal(again.bits, host.bits)
        assert iters[1] < 400
