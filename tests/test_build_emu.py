"""The GPU stream producer's algorithm (csrc/kd_build.cu), run on the CPU by tests/emu/emu_build.cpp with the same
element functions and the same parallel formulations, must give the byte-identical stream of the host compressor
(kd_compress.cpp, itself pinned to the unmodified reference build() in test_host_codec.py)."""
import ctypes

import numpy as np
import pytest

import oracle_api as oa
from volumerenderer_b200 import synth

pytestmark = pytest.mark.skipif(oa.EMU is None or not hasattr(oa.EMU, "emu_build"), reason="tests/emu/libemu.so not built")


def _brick(kind, X, Y, Z):
    if kind == "noise":
        return np.random.RandomState(11).randint(0, 256, (Z, Y, X)).astype(np.uint8)
    if kind == "constant":
        return np.full((Z, Y, X), 77, np.uint8)
    if kind == "ramp":
        z, y, x = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
        return ((x * 3 + y * 2 + z * 5) % 256).astype(np.uint8)
    # a scaled-down mixing layer: empty below, constant above, noisy interface
    z, y, x = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    h = Z / 2 + (Z / 6) * np.sin(2 * np.pi * x / X) * np.cos(2 * np.pi * y / Y)
    d = (z - h) / max(Z / 10, 1)
    rng = np.random.RandomState(7)
    v = np.where(d <= -1, 0, np.where(d >= 1, 230, np.clip(np.round(115 * (1 + d)) + rng.randint(-16, 17, d.shape) * (1 - np.abs(d)), 0, 255)))
    return v.astype(np.uint8)


@pytest.mark.parametrize("kind", ["rm", "noise", "ramp", "constant"])
@pytest.mark.parametrize("shape", [(32, 32, 16), (64, 64, 32), (16, 8, 4)])
@pytest.mark.parametrize("tol,epochs", [(1, 2), (6, 1), (0, 5), (4, 2)])
def test_parallel_formulation_matches_host(kind, shape, tol, epochs):
    X, Y, Z = shape
    vol = np.ascontiguousarray(_brick(kind, X, Y, Z))
    h = ctypes.c_void_p()
    it = (ctypes.c_longlong * 2)()
    assert oa.EMU.emu_build(vol.ctypes.data, X, Y, Z, tol, epochs, ctypes.byref(h), it) == 0
    assert oa.EMU.emu_build_matches_host(h, vol.ctypes.data, X, Y, Z, tol, epochs) == 1
    assert it[1] < 400, f"running-mean fixed point needed {it[1]} passes on one level"


def test_order_exact_accumulation():
    rng = np.random.RandomState(3)
    for n, hi, x0 in [(1, 10, 0.0), (1000, 65026, 0.3), (70000, 65026, 123.456789), (300000, 5, 1e-9), (200000, 65026, 64999.999),
                      (100000, 2, 0.0), (50000, 65026, 2.0 ** 30 + 0.5), (123457, 1000, 7.0 / 3.0)]:
        e = rng.randint(0, hi, n).astype(np.uint32)
        e[rng.rand(n) < 0.3] = 0
        assert oa.EMU.emu_accumulate_check(x0, e.ctypes.data, n) == 1, (n, hi, x0)
    z = np.zeros(100000, np.uint32)
    z[77777] = 9
    assert oa.EMU.emu_accumulate_check(0.1, z.ctypes.data, z.size) == 1
