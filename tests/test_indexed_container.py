"""CPU: the indexed ("random-access") container of csrc/kd_kdx.cpp: deterministic bytes, opening without a walk
gives the index of vr_stream_prepare and the stream of VolumeKdtree::save bit for bit, tile subsets fetch only their byte
ranges from disk, damaged files are refused."""
import glob
import os

import numpy as np
import pytest

import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


def _same_index(a, b):
    ia, ib = a.index_arrays(), b.index_arrays()
    for k in ia:
        assert np.array_equal(ia[k], ib[k]), k
    assert a.index_info() == b.index_info() or {**a.index_info(), "subset": False} == {**b.index_info(), "subset": False}


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_roundtrip_on_reference_goldens(path, tmp_path):
    g = np.load(path)
    raw = tmp_path / "tree.bin"
    raw.write_bytes(g["container"].tobytes())
    s = vr.Stream.open(raw)            # the reference's own container
    f = tmp_path / "tree.kdx"
    s.save_indexed(f)
    blob = np.fromfile(f, np.uint8)
    assert np.array_equal(blob, s.serialize_indexed())            # file == serialised image
    t = vr.Stream.open_indexed(f, verify=True)
    assert bytes(t.header) == bytes(s.header)
    assert np.array_equal(t.distance_map, s.distance_map)
    assert np.array_equal(t.bits, s.bits)
    _same_index(s, t)
    assert blob.size - 16 * 9 <= t.bytes_read <= blob.size             # everything but alignment padding
    # the plain container written back from the indexed one is the reference's file again
    back = tmp_path / "back.bin"
    t.save(back)
    assert back.read_bytes() == g["container"].tobytes()
    # deterministic: a second stream object made from the same bytes serialises to the same image
    u = vr.Stream.from_memory(s.header, s.distance_map, s.bits)
    assert np.array_equal(u.serialize_indexed(), blob)
    v = vr.Stream.from_memory_indexed(blob, verify=True)
    assert np.array_equal(v.serialize_indexed(), blob)


def test_tile_subset_reads_only_its_byte_ranges(tmp_path):
    vol = synth.shell_brick((64, 64, 64), noise=9, seed=3)         # 16 tiles of 32x32x16
    s, _ = vr.Stream.compress(vol, 1, 2)
    f = tmp_path / "b.kdx"
    s.save_indexed(f)
    info = s.index_info()
    ix = s.index_arrays()
    assert info["tiles"] == 16 and info["live_tiles"] > 4
    live = np.flatnonzero(ix["tileLeafBlock"] != 0xFFFFFFFF)
    pick = live[[1, len(live) // 2]]
    mask = np.zeros(info["tiles"], np.uint8)
    mask[pick] = 1
    t = vr.Stream.open_indexed(f, tile_mask=mask)
    assert t.index_info()["subset"]
    _same_index(s, t)
    full, part = s.bits, t.bits
    want = 0
    for tile in pick:
        b = ix["tileLeafBlock"][tile]
        lo, hi = ix["blockSymBegin"][b] // 4, (int(ix["blockSymEnd"][b]) + 3) // 4
        assert np.array_equal(part[lo:hi], full[lo:hi])
        want += hi - lo
    assert np.count_nonzero(part) <= want
    header_and_index = os.path.getsize(f) - ((full.size + 15) // 16) * 16
    assert t.bytes_read <= header_and_index + want
    assert t.bytes_read < os.path.getsize(f)
    # a subset is not the stream: it cannot be saved, validated or re-serialised
    for call in (lambda: t.save(tmp_path / "x.bin"), t.validate, t.serialize_indexed):
        with pytest.raises(_lib.VrError) as e:
            call()
        assert e.value.code == _lib.VR_ERR_STATE


def test_damaged_files_are_refused(tmp_path):
    vol = synth.shell_brick((32, 32, 32), noise=5, seed=1)
    s, _ = vr.Stream.compress(vol, 1, 2)
    blob = s.serialize_indexed()

    def opens(b, verify=False):
        try:
            vr.Stream.from_memory_indexed(b, verify=verify)
            return None
        except _lib.VrError as e:
            return e.code

    assert opens(blob, True) is None
    bad = blob.copy(); bad[0] ^= 1                                   # magic
    assert opens(bad) == _lib.VR_ERR_FORMAT
    assert opens(blob[: blob.size - 16]) == _lib.VR_ERR_IO           # truncated
    bad = blob.copy(); bad[-20] ^= 0x40                              # a stream byte: only the checksum can tell
    assert opens(bad, False) is None and opens(bad, True) == _lib.VR_ERR_FORMAT
    # a leaf offset outside its tile's symbol range
    import struct
    secs = [struct.unpack_from("<QQ", blob, 352 - 16 * 9 + 16 * k) for k in range(9)]
    off7, n7 = secs[7]
    leaf = blob[off7:off7 + n7].view(np.uint32)
    k = int(np.flatnonzero(leaf != 0xFFFFFFFF)[3])
    bad = blob.copy(); bad[off7 + 4 * k: off7 + 4 * k + 4] = np.frombuffer(struct.pack("<I", s.header.numActiveNodes + 5), np.uint8)
    assert opens(bad) == _lib.VR_ERR_FORMAT
    with pytest.raises(_lib.VrError) as e:
        vr.Stream.open_indexed(tmp_path / "missing.kdx")
    assert e.value.code == _lib.VR_ERR_IO
