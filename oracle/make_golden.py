"""TEST INFRASTRUCTURE.  Generates tests/golden/*.npz by running the UNMODIFIED reference
(oracle/_ref/libvr_ref.so, built by oracle/Makefile from /root/reference) on small seeded inputs:
input volume, the reference's stream (container bytes as VolumeKdtree::save writes them, checked
against an actual save() call) and the reference's levelCut output.  Run in the build container
only (the GPU box has no /root/reference); the fixtures are committed.
    python oracle/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_api as oa  # noqa: E402
from volumerenderer_b200 import synth  # noqa: E402

CASES = [
    ("shell_16x16x8_t1e2", lambda: synth.shell_brick((16, 16, 8)), 1, 2),
    ("shellnoise_32x32x16_t1e2", lambda: synth.shell_brick((32, 32, 16), noise=8, seed=3), 1, 2),
    ("shellnoise_32x32x16_t6e5", lambda: synth.shell_brick((32, 32, 16), noise=8, seed=3), 6, 5),
    ("noise_8x8x8_t0e5", lambda: synth.noise_brick((8, 8, 8)), 0, 5),
    ("noise_16x16x16_t2e1", lambda: synth.noise_brick((16, 16, 16), seed=7), 2, 1),
    ("const_16x8x4_t1e2", lambda: synth.constant_brick((16, 8, 4)), 1, 2),
    ("rm_32x32x32_t1e2", lambda: synth.rm_region(512, 1024, 944, 32, 32, 32), 1, 2),
    ("slab_64x2x4_t1e2", lambda: synth.rm_region(100, 7, 950, 64, 2, 4), 1, 2),
    ("voxel_1x1x1_t1e2", lambda: np.full((1, 1, 1), 200, np.uint8), 1, 2),
]


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, gen, tol, ep in CASES:
        vol = gen()
        s = oa.ref_build(vol, tol, ep)
        dec = oa.ref_levelcut(s)
        blob = oa.container_bytes(s["Dmax"], s["D0"], s["X"], s["Y"], s["Z"], s["n"], s["dmap"], s["bits"])
        with tempfile.TemporaryDirectory() as td:
            p = os.path.join(td, "t.bin")
            oa.REF.ref_kd_save(s["h"], p.encode())
            assert open(p, "rb").read() == blob, "container_bytes() disagrees with the reference's save()"
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), volume=vol, container=np.frombuffer(blob, np.uint8),
                            decoded=dec, tolerance=tol, maxEpochs=ep)
        print(name, "nodes", s["n"], "maxerr", int(np.abs(dec.astype(int) - vol.astype(int)).max()))
        oa.REF.ref_kd_free(s["h"])


MRT_CASES = [
    ("mrt_shell_32x32x16_t1e2", lambda: synth.shell_brick((32, 32, 16), noise=4), 1, 2),
    ("mrt_odd_16x8x4_t1e2", lambda: synth.shell_brick((16, 8, 4), noise=9), 1, 2),
]


def main_mrt():
    """tests/golden/mrt/*.npz: the file MidRangeTree::save of the unmodified reference writes, the class's public arrays and
    MidRangeTree::levelCut's output."""
    out_dir = os.path.join(ROOT, "tests", "golden", "mrt")
    os.makedirs(out_dir, exist_ok=True)
    LL = oa.c_longlong
    for name, gen, tol, ep in MRT_CASES:
        vol = gen()
        Z, Y, X = vol.shape
        h = oa.REF.ref_mrt_build(vol.ctypes.data, X, Y, Z, tol, ep, 0)
        m = (LL * 7)()
        oa.REF.ref_mrt_meta(h, m)
        dm = np.zeros(m[0] + 1, np.uint8); dmr = np.zeros(m[0] + 1, np.uint8)
        bits = np.zeros(m[6], np.uint8); bitsr = np.zeros(m[6], np.uint8)
        oa.REF.ref_mrt_get(h, dm.ctypes.data, dmr.ctypes.data, bits.ctypes.data, bitsr.ctypes.data)
        dec = np.zeros((Z, Y, X), np.uint8)
        oa.REF.ref_mrt_levelcut(h, dec.ctypes.data)
        with tempfile.TemporaryDirectory() as td:
            p = os.path.join(td, "t.mrt")
            oa.REF.ref_mrt_save(h, p.encode())
            blob = np.fromfile(p, np.uint8)
        oa.REF.ref_mrt_free(h)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), volume=vol, container=blob, dmap=dm, dmap_range=dmr, bits=bits, bits_range=bitsr,
                            decoded=dec, meta=np.array(list(m), np.int64))
        print(name, "nodes", m[5], "file bytes", blob.size)


if __name__ == "__main__":
    if "--mrt-only" not in sys.argv:
        main()
    main_mrt()
