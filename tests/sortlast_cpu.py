"""TEST INFRASTRUCTURE: CPU stand-in for volumerenderer_b200.sortlast.CudaSlabOps, backed by the oracle's
sort-last restatement (oracle/render_oracle.cpp).  Lets the host-side exchange logic of
volumerenderer_b200/sortlast.py run on CPU tensors (lock-step in one process, or world_size-2 gloo)."""
import ctypes
from ctypes import c_int, c_void_p

import numpy as np
import torch

import oracle_api as oa
from volumerenderer_b200 import render

O = oa.ORACLE
O.oracle_partial_composite.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int]
O.oracle_partial_iso.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int]
O.oracle_resolve_composite.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int]
O.oracle_finish_composite.argtypes = [c_void_p, c_void_p, c_void_p, c_int]
for f in (O.oracle_partial_composite, O.oracle_partial_iso, O.oracle_resolve_composite, O.oracle_finish_composite):
    f.restype = None


class OracleSlabOps:
    def __init__(self, sortlast, desc, slab, width, height, threads=4):
        """slab: uint8 numpy (SZ, SY, SX) = the stored box of this rank (halo included)."""
        self.sl, self.slab, self.threads = sortlast, np.ascontiguousarray(slab), threads
        SZ, SY, SX = self.slab.shape
        self.geom = np.array(list(desc.global_) + list(desc.origin) + [SX, SY, SZ] + list(desc.own_lo) + list(desc.own_hi), np.int32)
        self.order = np.array([sortlast.order(o) for o in range(8)], np.int32)
        self.W, self.H = width, height

    def _blob(self, p):
        b = render.oracle_param_blob(p)
        return ctypes.create_string_buffer(b, len(b))

    def partial_composite(self, p):
        cseg = np.zeros((self.H, self.W), np.float32); tseg = np.zeros((self.H, self.W), np.float32)
        O.oracle_partial_composite(self.slab.ctypes.data, self.geom.ctypes.data, self._blob(p), cseg.ctypes.data, tseg.ctypes.data, self.threads)
        return torch.from_numpy(cseg), torch.from_numpy(tseg)

    def resolve_composite(self, p, tseg_all, cseg):
        ta = np.ascontiguousarray(tseg_all.numpy()); cs = np.ascontiguousarray(cseg.numpy())
        out = np.zeros((self.H, self.W, 2), np.float32)
        O.oracle_resolve_composite(self.slab.ctypes.data, self.geom.ctypes.data, self._blob(p), ta.ctypes.data, self.sl.n, self.sl.rank,
                                   self.order.ctypes.data, cs.ctypes.data, out.ctypes.data, self.threads)
        return torch.from_numpy(out)

    def finish_composite(self, p, total):
        t = np.ascontiguousarray(total.numpy()); out = np.zeros((self.H, self.W, 4), np.float32)
        O.oracle_finish_composite(self._blob(p), t.ctypes.data, out.ctypes.data, self.threads)
        return torch.from_numpy(out)

    def partial_iso(self, p):
        rgba = np.zeros((self.H, self.W, 4), np.float32); hit = np.zeros((self.H, self.W), np.int32)
        O.oracle_partial_iso(self.slab.ctypes.data, self.geom.ctypes.data, self._blob(p), rgba.ctypes.data, hit.ctypes.data, self.threads)
        return torch.from_numpy(rgba), torch.from_numpy(hit)

    def resolve_iso(self, p, hit, hit_min, rgba):
        none = hit_min == 0x7FFFFFFF
        rgba[(hit != hit_min) & ~none] = 0.0
        rgba[none] = 1.0 if self.sl.rank == 0 else 0.0
        return rgba

    def pack_box(self, origin, extent):
        x, y, z = origin; ex, ey, ez = extent
        return torch.from_numpy(np.ascontiguousarray(self.slab[z:z + ez, y:y + ey, x:x + ex]))

    def unpack_box(self, origin, extent, buf):
        x, y, z = origin; ex, ey, ez = extent
        self.slab[z:z + ez, y:y + ey, x:x + ex] = buf.numpy()

    def empty_box(self, extent):
        return torch.empty((extent[2], extent[1], extent[0]), dtype=torch.uint8)


def cut_slab(full, desc, stored, with_halo=True):
    """The stored box of a rank cut out of the whole (Z, Y, X) volume; with_halo=False leaves the halo
    zero-filled (to be completed by the halo exchange)."""
    (o, e) = stored
    box = np.ascontiguousarray(full[o[2]:o[2] + e[2], o[1]:o[1] + e[1], o[0]:o[0] + e[0]]).copy()
    if not with_halo:
        keep = np.zeros_like(box, bool)
        lo = [desc.own_lo[a] - o[a] for a in range(3)]; hi = [desc.own_hi[a] - o[a] for a in range(3)]
        keep[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]] = True
        box[~keep] = 0
    return box
