"""CPU: the oracle's C restatement of levelCut against (1) committed golden vectors produced by the
unmodified reference and (2) the reference itself when oracle/_ref is present."""
import glob
import os

import numpy as np
import pytest

import oracle_api as oa
from volumerenderer_b200 import synth

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _parse(blob):
    meta = (oa.c_int64 * 8)()
    buf = np.frombuffer(blob, np.uint8)
    assert oa.ORACLE.oracle_parse_container(buf.ctypes.data, buf.size, meta) == 0
    Dmax, D0, X, Y, Z, n, od, ob = list(meta)
    return Dmax, D0, X, Y, Z, n, buf[od:od + Dmax + 1], buf[ob:ob + (n + 3) // 4]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_levelcut_matches_golden(path):
    g = np.load(path)
    Dmax, D0, X, Y, Z, n, dm, bits = _parse(g["container"].tobytes())
    assert (Z, Y, X) == g["volume"].shape
    rc, out = oa.oracle_levelcut(bits, n, dm, D0, Dmax, X, Y, Z)
    assert rc == 0
    assert np.array_equal(out, g["decoded"])


def test_golden_present():
    assert len(GOLDEN) >= 9


@pytest.mark.skipif(oa.REF is None, reason="oracle/_ref not built (no reference sources)")
@pytest.mark.parametrize("dims", [(2, 1, 1), (4, 4, 2), (16, 16, 8), (32, 32, 16), (16, 4, 64), (128, 2, 2)])
@pytest.mark.parametrize("tol,ep", [(1, 2), (6, 5), (0, 5)])
def test_oracle_levelcut_matches_reference(dims, tol, ep):
    for vol in (synth.shell_brick(dims, noise=5, seed=tol), synth.noise_brick(dims, seed=ep)):
        s = oa.ref_build(vol, tol, ep)
        rc, out = oa.oracle_levelcut(s["bits"], s["n"], s["dmap"], s["D0"], s["Dmax"], s["X"], s["Y"], s["Z"])
        assert rc == 0
        assert np.array_equal(out, oa.ref_levelcut(s))
        oa.REF.ref_kd_free(s["h"])


def test_oracle_rejects_truncated_stream():
    g = np.load([p for p in GOLDEN if "shell_16x16x8" in p][0])
    Dmax, D0, X, Y, Z, n, dm, bits = _parse(g["container"].tobytes())
    rc, _ = oa.oracle_levelcut(bits, n - 5, dm, D0, Dmax, X, Y, Z)
    assert rc != 0


@pytest.mark.skipif(oa.REF is None, reason="oracle/_ref not built")
def test_oracle_nibble_pack_matches_midrangetree():
    vol = synth.shell_brick((32, 32, 16), noise=4)
    Z, Y, X = vol.shape
    h = oa.REF.ref_mrt_build(vol.ctypes.data, X, Y, Z, 1, 2, 0)
    m = (oa.c_longlong * 7)()
    oa.REF.ref_mrt_meta(h, m)
    Dmax, n, nb = m[0], m[5], m[6]
    dm = np.zeros(Dmax + 1, np.uint8); dmr = np.zeros(Dmax + 1, np.uint8)
    bits = np.zeros(nb, np.uint8); bitsr = np.zeros(nb, np.uint8)
    oa.REF.ref_mrt_get(h, dm.ctypes.data, dmr.ctypes.data, bits.ctypes.data, bitsr.ctypes.data)
    L = oa.REF.ref_mrt_bytearray(h, None)
    want = np.zeros(L, np.uint8)
    oa.REF.ref_mrt_bytearray(h, want.ctypes.data)
    assert oa.ORACLE.oracle_pack_nibbles(bits.ctypes.data, bitsr.ctypes.data, n, None) == L
    got = np.zeros(L, np.uint8)
    oa.ORACLE.oracle_pack_nibbles(bits.ctypes.data, bitsr.ctypes.data, n, got.ctypes.data)
    assert np.array_equal(got, want)
    # the midpoint tree of a MidRangeTree decodes with the same grammar
    rc, out = oa.oracle_levelcut(bits, n, dm, m[1], Dmax, X, Y, Z)
    ref = np.zeros((Z, Y, X), np.uint8)
    oa.REF.ref_mrt_levelcut(h, ref.ctypes.data)
    assert rc == 0 and np.array_equal(out, ref)
    oa.REF.ref_mrt_free(h)
