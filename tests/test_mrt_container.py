"""MidRangeTree::save / open container (MidRangeTree.cpp:753-833) through the product ABI, against files written by the
UNMODIFIED reference (tests/golden/mrt/*.npz from oracle/make_golden.py): vr_mrt_open returns the class's arrays, vr_mrt_save
writes the reference's bytes, and (-m gpu) the opened midpoint stream decodes to MidRangeTree::levelCut's output."""
import glob
import os

import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "mrt", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


def _header(g):
    m = g["meta"]
    hd = _lib.StreamHeader()
    Dmax, D0, X, Y, Z, n = [int(v) for v in m[:6]]
    hd.rootMax[:] = [X, Y, Z]
    hd.maxTreeDepth, hd.origTreeDepth, hd.X, hd.Y, hd.Z, hd.numActiveNodes = Dmax, D0, X, Y, Z, n
    return hd


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_open_and_save_match_the_reference_file(path, tmp_path):
    assert GOLDEN
    g = np.load(path)
    f = tmp_path / "ref.mrt"
    f.write_bytes(g["container"].tobytes())
    s, dmr, rb = vr.mrt_open(f)
    hd = s.header
    n = int(g["meta"][5])
    assert (hd.maxTreeDepth, hd.origTreeDepth, hd.X, hd.Y, hd.Z, hd.numActiveNodes) == tuple(int(v) for v in g["meta"][:6])
    assert np.array_equal(s.distance_map, g["dmap"]) and np.array_equal(dmr, g["dmap_range"])
    nb = (n + 3) // 4
    assert np.array_equal(s.bits[:nb - 1], g["bits"][:nb - 1])  # the last byte's unused symbols are normalised to 0
    assert np.array_equal(rb, g["bits_range"])                  # range stream byte for byte (the reference's open shifts it by 4)
    assert s.validate() == n
    # the writer: same bytes as the reference's save()
    out = tmp_path / "ours.mrt"
    vr.mrt_save(out, _header(g), g["dmap"], g["dmap_range"], g["bits"], g["bits_range"])
    assert out.read_bytes() == g["container"].tobytes()
    # the 4-bit shader stream from the opened file == from the class arrays
    a = vr.nibbles_pack(s.bits, rb, n)
    b = vr.nibbles_pack(g["bits"], g["bits_range"], n)
    assert np.array_equal(a, b)
    # damaged files
    (tmp_path / "short.mrt").write_bytes(g["container"].tobytes()[:100])
    with pytest.raises(vr.VrError):
        vr.mrt_open(tmp_path / "short.mrt")
    with pytest.raises(vr.VrError) as e:
        vr.mrt_open(tmp_path / "none.mrt")
    assert e.value.code == _lib.VR_ERR_IO


@pytest.mark.skipif(oa.REF is None or not hasattr(oa.REF, "ref_mrt_save"), reason="oracle/_ref not built")
def test_live_reference_save(tmp_path):
    from volumerenderer_b200 import synth
    vol = synth.rm_region(700, 1100, 940, 32, 32, 32)
    Z, Y, X = vol.shape
    h = oa.REF.ref_mrt_build(vol.ctypes.data, X, Y, Z, 1, 2, 0)
    f = tmp_path / "live.mrt"
    oa.REF.ref_mrt_save(h, str(f).encode())
    m = (oa.c_longlong * 7)()
    oa.REF.ref_mrt_meta(h, m)
    dm = np.zeros(m[0] + 1, np.uint8); dmr = np.zeros(m[0] + 1, np.uint8)
    bits = np.zeros(m[6], np.uint8); bitsr = np.zeros(m[6], np.uint8)
    oa.REF.ref_mrt_get(h, dm.ctypes.data, dmr.ctypes.data, bits.ctypes.data, bitsr.ctypes.data)
    oa.REF.ref_mrt_free(h)
    s, dmr2, rb = vr.mrt_open(f)
    assert np.array_equal(dmr2, dmr) and np.array_equal(rb, bitsr) and np.array_equal(s.distance_map, dm)
    out = tmp_path / "ours.mrt"
    vr.mrt_save(out, s.header, dm, dmr, bits, bitsr)
    assert out.read_bytes() == f.read_bytes()


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_gpu_decode_of_an_opened_mrt_file(path, tmp_path):
    g = np.load(path)
    f = tmp_path / "ref.mrt"
    f.write_bytes(g["container"].tobytes())
    s, _, _ = vr.mrt_open(f)
    tree = vr.VolumeKdtree()
    tree._stream = s
    hd = s.header
    tree.X, tree.Y, tree.Z = hd.X, hd.Y, hd.Z
    got = tree.levelCut(hd.maxTreeDepth)
    assert np.array_equal(got, g["decoded"])
