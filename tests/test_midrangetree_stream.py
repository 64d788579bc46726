"""MidRangeTree's 4-bit/node shader stream (SURVEY a-5) through the product ABI: vr_nibbles_pack equals the
unmodified reference's convertToByteArray byte for byte, unpack inverts it, and a stream rebuilt from the
4-bit form decodes on the GPU (-m gpu) to exactly MidRangeTree::levelCut's output."""
import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

pytestmark = pytest.mark.skipif(oa.REF is None, reason="oracle/_ref (unmodified reference build) not present")


def ref_mrt(vol, tol=1, epochs=2):
    Z, Y, X = vol.shape
    h = oa.REF.ref_mrt_build(vol.ctypes.data, X, Y, Z, tol, epochs, 0)
    m = (oa.c_longlong * 7)()
    oa.REF.ref_mrt_meta(h, m)
    Dmax, D0, n, nb = m[0], m[1], m[5], m[6]
    dm = np.zeros(Dmax + 1, np.uint8); dmr = np.zeros(Dmax + 1, np.uint8)
    bits = np.zeros(nb, np.uint8); bitsr = np.zeros(nb, np.uint8)
    oa.REF.ref_mrt_get(h, dm.ctypes.data, dmr.ctypes.data, bits.ctypes.data, bitsr.ctypes.data)
    L = oa.REF.ref_mrt_bytearray(h, None)
    packed = np.zeros(L, np.uint8)
    oa.REF.ref_mrt_bytearray(h, packed.ctypes.data)
    dec = np.zeros((Z, Y, X), np.uint8)
    oa.REF.ref_mrt_levelcut(h, dec.ctypes.data)
    oa.REF.ref_mrt_free(h)
    hd = _lib.StreamHeader()
    hd.rootMax[:] = [X, Y, Z]
    hd.maxTreeDepth, hd.origTreeDepth, hd.X, hd.Y, hd.Z, hd.numActiveNodes = Dmax, D0, X, Y, Z, n
    return dict(hd=hd, n=n, dm=dm, bits=bits, bitsr=bitsr, packed=packed, dec=dec)


CASES = [("shell", lambda: synth.shell_brick((32, 32, 16), noise=4)), ("rm", lambda: synth.rm_region(700, 1100, 940, 32, 32, 32)),
         ("odd", lambda: synth.shell_brick((16, 8, 4), noise=9))]


@pytest.mark.parametrize("name,make", CASES)
def test_pack_unpack_match_reference(name, make):
    r = ref_mrt(make())
    got = vr.nibbles_pack(r["bits"], r["bitsr"], r["n"])
    assert got.size == r["packed"].size and (got.size & (got.size - 1)) == 0  # power-of-two length
    assert np.array_equal(got, r["packed"])
    nb = (r["n"] + 3) // 4
    t, g = vr.nibbles_unpack(got, r["n"])
    mask = np.full(nb, 0xFF, np.uint8)
    if r["n"] % 4:
        mask[-1] = (1 << (2 * (r["n"] % 4))) - 1
    assert np.array_equal(t, r["bits"][:nb] & mask) and np.array_equal(g, r["bitsr"][:nb] & mask)
    s, rng = vr.Stream.from_nibbles(r["hd"], r["dm"], got)
    assert s.validate() == r["n"] and np.array_equal(rng, g)
    with pytest.raises(vr.VrError):
        vr.Stream.from_nibbles(r["hd"], r["dm"], got[: r["n"] // 2 - 1])  # truncated stream


@pytest.mark.gpu
@pytest.mark.parametrize("name,make", CASES)
def test_gpu_decode_from_4bit_stream(name, make):
    r = ref_mrt(make())
    s, _ = vr.Stream.from_nibbles(r["hd"], r["dm"], r["packed"])
    tree = vr.VolumeKdtree()
    tree._stream = s
    tree.X, tree.Y, tree.Z = r["hd"].X, r["hd"].Y, r["hd"].Z
    got = tree.levelCut(r["hd"].maxTreeDepth)
    assert np.array_equal(got, r["dec"])
