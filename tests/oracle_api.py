"""TEST INFRASTRUCTURE: ctypes access to oracle/liboracle.so (C restatements), oracle/_ref/libvr_ref.so
(the unmodified reference compiled by oracle/Makefile) and tests/emu/libemu.so (CPU emulation of the
device decode logic)."""
import ctypes
import os
import struct
from ctypes import c_int, c_int64, c_longlong, c_void_p

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = ctypes.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
_refp = os.path.join(ROOT, "oracle", "_ref", "libvr_ref.so")
REF = ctypes.CDLL(_refp) if os.path.exists(_refp) else None
_emup = os.path.join(ROOT, "tests", "emu", "libemu.so")
EMU = ctypes.CDLL(_emup) if os.path.exists(_emup) else None

ORACLE.oracle_levelcut.restype = c_int
ORACLE.oracle_levelcut.argtypes = [c_void_p, c_int64, c_void_p, c_int, c_int, c_int64, c_int64, c_int64, c_void_p]
ORACLE.oracle_pack_nibbles.restype = c_int64
ORACLE.oracle_pack_nibbles.argtypes = [c_void_p, c_void_p, c_int64, c_void_p]
ORACLE.oracle_parse_container.restype = c_int
ORACLE.oracle_parse_container.argtypes = [c_void_p, c_int64, c_void_p]
ORACLE.oracle_render.restype = c_int64
ORACLE.oracle_render.argtypes = [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int]
ORACLE.oracle_render_rows.restype = c_int64
ORACLE.oracle_render_rows.argtypes = [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int]

if REF is not None:
    LL = c_longlong
    REF.ref_kd_build.restype = c_void_p
    REF.ref_kd_build.argtypes = [c_void_p, LL, LL, LL, c_int, c_int, c_int]
    REF.ref_kd_open.restype = c_void_p
    REF.ref_kd_open.argtypes = [ctypes.c_char_p]
    REF.ref_kd_save.argtypes = [c_void_p, ctypes.c_char_p]
    REF.ref_kd_free.argtypes = [c_void_p]
    REF.ref_kd_meta.argtypes = [c_void_p, c_void_p]
    REF.ref_kd_get.argtypes = [c_void_p, c_void_p, c_void_p]
    REF.ref_kd_levelcut.argtypes = [c_void_p, c_void_p]
    REF.ref_kd_levelcut_bench.restype = ctypes.c_double
    REF.ref_kd_levelcut_bench.argtypes = [c_void_p, c_int, c_int, c_int]
    REF.ref_mrt_build.restype = c_void_p
    REF.ref_mrt_build.argtypes = [c_void_p, LL, LL, LL, c_int, c_int, c_int]
    REF.ref_mrt_free.argtypes = [c_void_p]
    REF.ref_mrt_meta.argtypes = [c_void_p, c_void_p]
    REF.ref_mrt_get.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    REF.ref_mrt_levelcut.argtypes = [c_void_p, c_void_p]
    REF.ref_mrt_bytearray.restype = LL
    REF.ref_mrt_bytearray.argtypes = [c_void_p, c_void_p]
    if hasattr(REF, "ref_mrt_save"):
        REF.ref_mrt_save.argtypes = [c_void_p, ctypes.c_char_p]

if EMU is not None:
    LL = c_longlong
    if hasattr(EMU, "emu_build"):
        EMU.emu_build.argtypes = [c_void_p, LL, LL, LL, c_int, c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(LL)]
        EMU.emu_build_matches_host.argtypes = [c_void_p, c_void_p, LL, LL, LL, c_int, c_int]
        EMU.emu_accumulate_check.argtypes = [ctypes.c_double, c_void_p, LL]
    EMU.emu_compress.restype = c_void_p
    EMU.emu_compress.argtypes = [c_void_p, LL, LL, LL, c_int, c_int, c_void_p]
    EMU.emu_from_memory.restype = c_void_p
    EMU.emu_from_memory.argtypes = [c_void_p, c_void_p, c_void_p, LL]
    EMU.emu_free.argtypes = [c_void_p]
    EMU.emu_meta.argtypes = [c_void_p, c_void_p]
    EMU.emu_get.argtypes = [c_void_p, c_void_p, c_void_p]
    EMU.emu_validate.argtypes = [c_void_p, c_void_p]
    EMU.emu_decode.argtypes = [c_void_p, c_void_p] + [LL] * 6 + [c_int]
    EMU.emu_decode6.argtypes = [c_void_p, c_void_p] + [LL] * 6 + [c_int]
    EMU.emu_decode6_stats.argtypes = [c_void_p]
    EMU.emu_check_chain_lut6.restype = LL
    EMU.emu_check_chain_lut6.argtypes = [ctypes.c_ulonglong]
    if hasattr(EMU, "emu_set_lazy"):
        EMU.emu_set_lazy.argtypes = [c_void_p, c_void_p, LL]
        EMU.emu_culled_count.restype = LL
    EMU.emu_check_chain_lut.restype = LL
    EMU.emu_check_chain_lut.argtypes = [ctypes.c_ulonglong]


def oracle_levelcut(bits, n, dmap, D0, Dmax, X, Y, Z):
    bits = np.ascontiguousarray(bits, np.uint8)
    dmap = np.ascontiguousarray(dmap, np.uint8)
    out = np.full((Z, Y, X), 0xCD, np.uint8)
    rc = ORACLE.oracle_levelcut(bits.ctypes.data, n, dmap.ctypes.data, D0, Dmax, X, Y, Z, out.ctypes.data)
    return rc, out


def ref_build(vol, tol, epochs, threads=0):
    """Runs the unmodified reference build(); returns dict(meta, dmap, bits, handle)."""
    vol = np.ascontiguousarray(vol, np.uint8)
    Z, Y, X = vol.shape
    h = REF.ref_kd_build(vol.ctypes.data, X, Y, Z, tol, epochs, threads)
    return ref_stream(h)


def ref_stream(h):
    m = (c_longlong * 7)()
    REF.ref_kd_meta(h, m)
    m = list(m)
    dm = np.zeros(m[0] + 1, np.uint8)
    bits = np.zeros(m[6], np.uint8)
    REF.ref_kd_get(h, dm.ctypes.data, bits.ctypes.data)
    return dict(Dmax=m[0], D0=m[1], X=m[2], Y=m[3], Z=m[4], n=m[5], dmap=dm, bits=bits, h=h)


def ref_levelcut(s):
    out = np.zeros((s["Z"], s["Y"], s["X"]), np.uint8)
    REF.ref_kd_levelcut(s["h"], out.ctypes.data)
    return out


def container_bytes(Dmax, D0, X, Y, Z, n, dmap, bits):
    """Serialises a stream exactly as VolumeKdtree::save does (VolumeKdTree_recover.cpp:535-544)."""
    hdr = struct.pack("<6q2i4q", 0, 0, 0, X, Y, Z, Dmax, D0, X, Y, Z, n)
    return hdr + bytes(np.asarray(dmap, np.uint8)[: Dmax + 1]) + bytes(np.asarray(bits, np.uint8)[: (n + 3) // 4])




def camera_params(eye, front, up, vfov_deg, W, H, step, iso, mode, znear=0.1, zfar=100.0):
    """GLM lookAt (RH) basis + perspectiveFov scale factors (main.cpp:396-397), packed like the oracle's Params."""
    eye = np.asarray(eye, np.float64); f = np.asarray(front, np.float64); up = np.asarray(up, np.float64)
    f = f / np.linalg.norm(f)
    s = np.cross(f, up); s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    th = 1.0 / np.tan(np.radians(vfov_deg) / 2.0)
    tw = th * H / W
    vals = list(eye) + list(s) + list(u) + list(f) + [tw, th, znear, zfar] + list(step) + [iso]
    return struct.pack("<20f3i", *[float(np.float32(v)) for v in vals], W, H, mode)


def oracle_render(vol, params, threads=8, want_hits=False):
    vol = np.ascontiguousarray(vol, np.uint8)
    Z, Y, X = vol.shape
    W, H, mode = struct.unpack_from("<3i", params, 80)
    out = np.zeros((H, W, 4), np.float32)
    hits = np.zeros((H, W), np.int32) if want_hits else None
    buf = ctypes.create_string_buffer(params, len(params))
    S = ORACLE.oracle_render(vol.ctypes.data, X, Y, Z, buf, out.ctypes.data,
                             hits.ctypes.data if want_hits else None, threads)
    return out, S, hits
