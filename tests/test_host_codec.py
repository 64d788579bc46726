"""CPU: host side of the product (no compute on a GPU): compressor vs the reference's build(),
container round trips, grammar validation, index-driven decode logic under CPU emulation, and the
C ABI symbol table."""
import ctypes
import glob
import os
import re

import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


def test_capi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vr_b200.h")).read()
    declared = set(re.findall(r"\b(vr_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(_lib.lib, name), f"libvr_b200.so does not export {name}"
    assert declared == set(_lib.EXPORTED), declared ^ set(_lib.EXPORTED)


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_compress_is_byte_identical_to_reference_golden(path):
    g = np.load(path)
    s, st = vr.Stream.compress(g["volume"], int(g["tolerance"]), int(g["maxEpochs"]))
    hd = s.header
    blob = oa.container_bytes(hd.maxTreeDepth, hd.origTreeDepth, hd.X, hd.Y, hd.Z, hd.numActiveNodes,
                              s.distance_map, s.bits)
    assert blob == g["container"].tobytes()
    assert st.zeroRunRewrites == 0  # SURVEY C-13
    assert s.validate() == hd.numActiveNodes


@pytest.mark.skipif(oa.REF is None, reason="oracle/_ref not built")
@pytest.mark.parametrize("dims", [(1, 1, 1), (2, 1, 1), (8, 8, 8), (32, 32, 16), (64, 64, 32), (16, 4, 64)])
def test_compress_matches_reference_live(dims):
    for kind, tol, ep in [("shell", 1, 2), ("noise", 6, 5), ("shell", 0, 5), ("noise", 2, 1), ("shell", 4, 3)]:
        vol = synth.shell_brick(dims, noise=7, seed=tol + ep) if kind == "shell" else synth.noise_brick(dims, seed=tol)
        r = oa.ref_build(vol, tol, ep)
        s, _ = vr.Stream.compress(vol, tol, ep)
        hd = s.header
        assert (hd.maxTreeDepth, hd.origTreeDepth, hd.numActiveNodes) == (r["Dmax"], r["D0"], r["n"])
        assert np.array_equal(s.distance_map, r["dmap"])
        assert np.array_equal(s.bits, r["bits"][: (r["n"] + 3) // 4])
        oa.REF.ref_kd_free(r["h"])


def test_container_roundtrip_and_open_slack(tmp_path):
    g = np.load(GOLDEN[1])
    p = tmp_path / "tree.bin"
    p.write_bytes(g["container"].tobytes())
    s = vr.Stream.open(p)
    q = tmp_path / "again.bin"
    s.save(q)
    assert q.read_bytes() == g["container"].tobytes()
    # the reference's open() over-sizes the bit array by 8 bytes (SURVEY C-3): both lengths load
    p2 = tmp_path / "slack.bin"
    p2.write_bytes(g["container"].tobytes() + b"\0" * 8)
    s2 = vr.Stream.open(p2)
    assert np.array_equal(s2.bits, s.bits)


@pytest.mark.skipif(oa.REF is None, reason="oracle/_ref not built")
def test_reference_opens_our_file_and_we_open_theirs(tmp_path):
    vol = synth.shell_brick((32, 32, 16), noise=3)
    s, _ = vr.Stream.compress(vol, 1, 2)
    p = tmp_path / "ours.bin"
    s.save(p)
    h = oa.REF.ref_kd_open(str(p).encode())
    r = oa.ref_stream(h)
    dec = oa.ref_levelcut(r)
    rc, want = oa.oracle_levelcut(s.bits, s.header.numActiveNodes, s.distance_map, s.header.origTreeDepth,
                                  s.header.maxTreeDepth, 32, 32, 16)
    assert rc == 0 and np.array_equal(dec, want)
    p2 = tmp_path / "theirs.bin"
    oa.REF.ref_kd_save(h, str(p2).encode())
    s2 = vr.Stream.open(p2)
    assert np.array_equal(s2.bits, s.bits) and s2.header.numActiveNodes == s.header.numActiveNodes
    oa.REF.ref_kd_free(h)


def test_error_codes_not_exits(tmp_path):
    with pytest.raises(vr.VrError) as e:
        vr.Stream.open(tmp_path / "missing.bin")
    assert e.value.code == _lib.VR_ERR_IO
    with pytest.raises(vr.VrError) as e:
        vr.Stream.compress(np.zeros((3, 4, 4), np.uint8), 1, 2)  # non power of two
    assert e.value.code == _lib.VR_ERR_FORMAT
    g = np.load([p for p in GOLDEN if "shell_16x16x8" in p][0])
    blob = bytearray(g["container"].tobytes())
    hd = _lib.StreamHeader.from_buffer_copy(bytes(blob[:88]))
    dm = np.frombuffer(bytes(blob[88:88 + hd.maxTreeDepth + 1]), np.uint8)
    bits = np.frombuffer(bytes(blob[88 + hd.maxTreeDepth + 1:]), np.uint8)
    with pytest.raises(vr.VrError):  # short bit array
        vr.Stream.from_memory(hd, dm, bits[:-3])
    hd.numActiveNodes -= 7  # truncated stream parses short
    s = vr.Stream.from_memory(hd, dm, bits)
    with pytest.raises(vr.VrError) as e:
        s.validate()
    assert e.value.code == _lib.VR_ERR_FORMAT


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_emulated_device_decode_matches_golden(path):
    """Runs kd_decode_core.h (the device parser + strip gather) thread by thread on the CPU."""
    g = np.load(path)
    blob = g["container"].tobytes()
    hd = _lib.StreamHeader.from_buffer_copy(blob[:88])
    dm = np.frombuffer(blob[88:88 + hd.maxTreeDepth + 1], np.uint8).copy()
    bits = np.frombuffer(blob[88 + hd.maxTreeDepth + 1:], np.uint8).copy()
    h = oa.EMU.emu_from_memory(ctypes.byref(hd), dm.ctypes.data, bits.ctypes.data, bits.size)
    assert h
    Z, Y, X = g["volume"].shape
    ox = 16 if X >= 16 else 3
    slab = np.full((Z + 3, Y + 2, X + 32), 0xAB, np.uint8)
    assert oa.EMU.emu_decode(h, slab.ctypes.data, X + 32, Y + 2, Z + 3, ox, 1, 2, 1 if X >= 16 else 0) == 0
    assert np.array_equal(slab[2:2 + Z, 1:1 + Y, ox:ox + X], g["decoded"])
    slab[2:2 + Z, 1:1 + Y, ox:ox + X] = 0xAB
    assert (slab == 0xAB).all(), "decode wrote outside its brick"
    oa.EMU.emu_free(h)


def test_chain_table_equals_sequential_chain_for_all_inputs():
    """16384 chains x 256 start values: the reference's distances 64..1 and two other sets whose sum
    fits the table's signed 8-bit offset (larger sums take the loop-based parser)."""
    for dist in ([64, 32, 16, 8, 4, 2, 1], [100, 0, 13, 7, 3, 3, 1], [1, 2, 4, 8, 16, 32, 64]):
        assert oa.EMU.emu_check_chain_lut(sum(d << (8 * k) for k, d in enumerate(dist))) == 0


def _full_size(kind):
    return {"rm": lambda: synth.rm_brick(3, 4, 7), "noise": synth.noise_brick,
            "shell": lambda: synth.shell_brick((256, 256, 128), noise=3)}[kind]()


@pytest.mark.parametrize("kind", ["rm", "noise", "shell"])
def test_emulated_fast_path_full_size_brick(kind):
    """256x256x128 (tiles of 2^14 voxels): the classic kernel's table-driven parser and its warp-unit gather."""
    vol = _full_size(kind)
    s, _ = vr.Stream.compress(vol, 1, 2)
    hd = s.header
    dm, bits = s.distance_map, s.bits
    h = oa.EMU.emu_from_memory(ctypes.byref(hd), dm.ctypes.data, bits.ctypes.data, bits.size)
    slab = np.full((128 + 2, 256 + 1, 256 + 32), 0xAB, np.uint8)
    assert oa.EMU.emu_decode(h, slab.ctypes.data, 256 + 32, 256 + 1, 128 + 2, 16, 1, 1, 1) == 0
    rc, want = oa.oracle_levelcut(bits, hd.numActiveNodes, dm, hd.origTreeDepth, hd.maxTreeDepth, 256, 256, 128)
    assert rc == 0
    assert np.array_equal(slab[1:129, 1:257, 16:272], want)
    slab[1:129, 1:257, 16:272] = 0xAB
    assert (slab == 0xAB).all()
    oa.EMU.emu_free(h)


def test_chain_table_of_the_persistent_kernel_for_all_inputs():
    for dist in ([64, 32, 16, 8, 4, 2, 1], [100, 0, 13, 7, 3, 3, 1], [1, 2, 4, 8, 16, 32, 64]):
        assert oa.EMU.emu_check_chain_lut6(sum(d << (8 * k) for k, d in enumerate(dist))) == 0


@pytest.mark.parametrize("kind,dims", [("rm", (256, 256, 128)), ("noise", (256, 256, 128)), ("shell", (256, 256, 128)),
                                       ("shell", (128, 128, 64)), ("shell", (256, 128, 128))])
def test_emulated_persistent_kernel_parser(kind, dims):
    """kd_parse6.h (the leaf parser of the persistent kernel: select words, chain table, static window re-alignment) and its
    work lists, lane by lane on the CPU, against the levelCut restatement, with the device's staging capacity of 304 chunks per
    unit (every unit must fit; too small a capacity is reported as a format error, never decoded wrongly).  256x128x128 is the
    other leaf-box orientation the kernel is instantiated for (last split along x instead of y)."""
    X, Y, Z = dims
    vol = _full_size(kind) if dims == (256, 256, 128) else synth.shell_brick(dims, noise=3)
    s, _ = vr.Stream.compress(vol, 1, 2)
    hd = s.header
    dm, bits = s.distance_map, s.bits
    h = oa.EMU.emu_from_memory(ctypes.byref(hd), dm.ctypes.data, bits.ctypes.data, bits.size)
    slab = np.full((Z + 2, Y + 1, X + 32), 0xAB, np.uint8)
    assert oa.EMU.emu_decode6(h, slab.ctypes.data, X + 32, Y + 1, Z + 2, 16, 1, 1, 304) == 0
    rc, want = oa.oracle_levelcut(bits, hd.numActiveNodes, dm, hd.origTreeDepth, hd.maxTreeDepth, X, Y, Z)
    assert rc == 0
    assert np.array_equal(slab[1:Z + 1, 1:Y + 1, 16:16 + X], want)
    slab[1:Z + 1, 1:Y + 1, 16:16 + X] = 0xAB
    assert (slab == 0xAB).all()
    st = (ctypes.c_longlong * 4)()
    oa.EMU.emu_decode6_stats(st)
    assert 0 < st[1] <= 304
    if kind == "noise":
        assert oa.EMU.emu_decode6(h, slab.ctypes.data, X + 32, Y + 1, Z + 2, 16, 1, 1, int(st[1]) - 1) == _lib.VR_ERR_FORMAT
    oa.EMU.emu_free(h)


@pytest.mark.parametrize("dims", [(64, 64, 32), (64, 32, 32)])
def test_longest_possible_units_fit_the_staging_area(dims):
    """A hand-made stream in which no node is pruned and every voxel carries the longest refinement chain (8 symbols per
    voxel item): the units of the persistent kernel are then as long as the grammar allows.  They must fit the staging
    capacity of 304 chunks (kd_decode.cu: at most 4 runs x 74 chunks), and the parser must decode them like the levelCut
    restatement.  Both leaf-box orientations."""
    X, Y, Z = dims
    s, _ = vr.Stream.compress(synth.shell_brick(dims, noise=3), 1, 2)  # header and distance map of a brick of this shape
    hd, dm = s.header, s.distance_map.copy()
    D0 = hd.origTreeDepth
    rng = np.random.default_rng(7)
    syms = np.empty(((1 << D0) - 1) + (8 << D0), np.uint8)
    pos, stack = 0, [0]
    while stack:  # preorder of the complete tree; symbols 0..2 only (3 would prune a node / end a chain)
        d = stack.pop()
        if d == D0:
            syms[pos:pos + 8] = rng.integers(0, 3, 8); pos += 8
        else:
            syms[pos] = rng.integers(0, 3); pos += 1
            stack += [d + 1, d + 1]
    assert pos == syms.size
    sy = np.concatenate([syms, np.zeros((-syms.size) % 4, np.uint8)]).reshape(-1, 4)
    packed = np.concatenate([(sy[:, 0] | (sy[:, 1] << 2) | (sy[:, 2] << 4) | (sy[:, 3] << 6)).astype(np.uint8), np.zeros(64, np.uint8)])
    hd.numActiveNodes = syms.size
    rc, want = oa.oracle_levelcut(packed, hd.numActiveNodes, dm, hd.origTreeDepth, hd.maxTreeDepth, X, Y, Z)
    assert rc == 0
    h = oa.EMU.emu_from_memory(ctypes.byref(hd), dm.ctypes.data, packed.ctypes.data, packed.size)
    slab = np.full((Z + 2, Y + 1, X + 32), 0xAB, np.uint8)
    assert oa.EMU.emu_decode6(h, slab.ctypes.data, X + 32, Y + 1, Z + 2, 16, 1, 1, 304) == 0
    assert np.array_equal(slab[1:Z + 1, 1:Y + 1, 16:16 + X], want)
    st = (ctypes.c_longlong * 4)()
    oa.EMU.emu_decode6_stats(st)
    assert 288 <= st[1] <= 296  # 18 431 symbols = 4 608 bytes per unit, plus the alignment of its 2 or 4 runs
    oa.EMU.emu_free(h)
