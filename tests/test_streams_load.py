"""CPU: vr_streams_load without a device context = VolumeKdtree::open (VolumeKdTree_recover.cpp:554-594) for a whole
timestep on worker threads -- same streams as one vr_stream_open per file, container bytes from the unmodified
reference (tests/golden), error behaviour on missing / malformed files."""
import glob
import os

import numpy as np
import pytest

import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz")))


def test_load_many_equals_open_each(tmp_path):
    paths = []
    for k in range(9):
        b = synth.rm_region(600 + 40 * k, 1000, 930, 64, 32, 32) if k % 3 else synth.constant_brick((64, 32, 32), 10 + k)
        s, _ = vr.Stream.compress(b, 1, 2)
        p = tmp_path / f"tree_{k}.bin"
        s.save(p)
        paths.append(p)
    many = vr.load_streams(paths, None, threads=4)
    assert len(many) == len(paths)
    for p, s in zip(paths, many):
        one = vr.Stream.open(p)
        assert bytes(one.header) == bytes(s.header)
        assert np.array_equal(one.bits, s.bits) and np.array_equal(one.distance_map, s.distance_map)
        assert s.validate() == s.header.numActiveNodes
        assert s.device_bytes()[0] > 0  # prepared: the host image exists


@pytest.mark.skipif(not GOLDEN, reason="no golden fixtures")
def test_load_reference_containers(tmp_path):
    paths, want = [], []
    for g in GOLDEN:
        z = np.load(g)
        p = tmp_path / (os.path.basename(g) + ".bin")
        p.write_bytes(bytes(z["container"]))
        paths.append(p)
        want.append(bytes(z["container"]))
    for s, w, p in zip(vr.load_streams(paths, None), want, paths):
        q = str(p) + ".resaved"
        s.save(q)
        assert open(q, "rb").read() == w  # byte-identical round trip of the reference's own files


def test_load_reports_missing_and_malformed(tmp_path):
    b = synth.shell_brick((32, 32, 16))
    s, _ = vr.Stream.compress(b, 1, 2)
    good = tmp_path / "good.bin"
    s.save(good)
    with pytest.raises(_lib.VrError) as ei:
        vr.load_streams([good, tmp_path / "missing.bin", good], None, threads=2)
    assert ei.value.code == _lib.VR_ERR_IO
    bad = tmp_path / "bad.bin"
    raw = bytearray(good.read_bytes())
    bad.write_bytes(bytes(raw[:len(raw) // 2]))  # truncated container
    with pytest.raises(_lib.VrError) as ei:
        vr.load_streams([good, bad], None, threads=2)
    assert ei.value.code in (_lib.VR_ERR_IO, _lib.VR_ERR_FORMAT)
