"""Host-side mirror of the reference's renderer set-up: the uniforms main.cpp uploads
(/root/reference/volume_renderer/main.cpp:319-341,396-402) and the shader-selectable mode
(main.cpp:71-74), on top of the C ABI's vr_render*."""
import ctypes
from ctypes import byref, c_float, c_int32, c_uint64, c_void_p

import numpy as np

from . import _lib
from ._lib import check, lib

COMPOSITE, ISOSURFACE, ISOSURFACE_COMPRESSED = 0, 1, 2


class RenderParams(ctypes.Structure):
    _fields_ = [("eye", c_float * 3), ("front", c_float * 3), ("up", c_float * 3), ("vfov_deg", c_float),
                ("znear", c_float), ("zfar", c_float), ("step_size", c_float * 3), ("isoValue", c_float),
                ("width", c_int32), ("height", c_int32), ("mode", c_int32)]


_lib.lib.vr_render_defaults.restype = ctypes.c_int
_lib.lib.vr_render_defaults.argtypes = [ctypes.POINTER(RenderParams)]
_lib.lib.vr_camera_basis.restype = ctypes.c_int
_lib.lib.vr_camera_basis.argtypes = [ctypes.POINTER(RenderParams), ctypes.POINTER(c_float)]
for _n in ("vr_render", "vr_render_to_host", "vr_render_to_host_static", "vr_render_compressed"):
    getattr(_lib.lib, _n).restype = ctypes.c_int
    getattr(_lib.lib, _n).argtypes = [c_void_p, c_void_p, ctypes.POINTER(RenderParams), c_void_p, c_void_p, c_void_p]
_lib.lib.vr_render_compressed_lazy.restype = ctypes.c_int
_lib.lib.vr_render_compressed_lazy.argtypes = [c_void_p, c_void_p, ctypes.POINTER(RenderParams), c_void_p, c_void_p, c_void_p,
                                               ctypes.POINTER(ctypes.c_int64)]
_lib.lib.vr_lazy_tile_mask.restype = ctypes.c_int
_lib.lib.vr_lazy_tile_mask.argtypes = [ctypes.POINTER(c_void_p), c_int32, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64),
                                       ctypes.POINTER(RenderParams), c_void_p, c_void_p, ctypes.POINTER(c_int32)]
_lib.lib.vr_stream_tile_bounds.restype = ctypes.c_int
_lib.lib.vr_stream_tile_bounds.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_int64, ctypes.POINTER(c_int32)]
_lib.lib.vr_static_scratch_max_pixels.restype = ctypes.c_int64
_lib.lib.vr_static_scratch_max_pixels.argtypes = []
_lib.EXPORTED[:] = sorted(set(_lib.EXPORTED) | {"vr_render_defaults", "vr_camera_basis", "vr_render", "vr_render_to_host_static",
                                                "vr_static_scratch_max_pixels",
                                                "vr_render_to_host", "vr_render_compressed", "vr_render_compressed_lazy",
                                                "vr_lazy_tile_mask", "vr_stream_tile_bounds"})


def default_params(width=1600, height=1200, mode=COMPOSITE, eye=None, front=None, up=None, iso=None, step=None):
    """Reference constants (main.cpp:27,33-40,52,330-331) with optional overrides."""
    p = RenderParams()
    check(lib.vr_render_defaults(byref(p)))
    p.width, p.height, p.mode = int(width), int(height), int(mode)
    if eye is not None:
        p.eye[:] = [float(v) for v in eye]
    if front is not None:
        p.front[:] = [float(v) for v in front]
    if up is not None:
        p.up[:] = [float(v) for v in up]
    if iso is not None:
        p.isoValue = float(iso)
    if step is not None:
        p.step_size[:] = [float(v) for v in step]
    return p


def oblique_params(width, height, mode=COMPOSITE, **kw):
    """Second view of SURVEY 8d: eye (-0.9, 0.6, -0.9) looking at the cube centre."""
    eye = np.array([-0.9, 0.6, -0.9])
    return default_params(width, height, mode, eye=eye, front=-eye / np.linalg.norm(eye), **kw)


def camera_basis(p):
    out = (c_float * 20)()
    check(lib.vr_camera_basis(byref(p), out))
    return np.array(out[:], np.float32)


def oracle_param_blob(p):
    """The same frame packed the way oracle/render_oracle.cpp expects (basis computed by the product's
    host code, so oracle and kernel see identical float inputs)."""
    b = camera_basis(p)
    return b.tobytes() + np.array([p.width, p.height, 0 if p.mode == COMPOSITE else 1], np.int32).tobytes()


def render_to_host(ctx, volume, p, want_hits=False, want_samples=True):
    """One frame -> (rgba float32 [H, W, 4] with row 0 at the bottom, samples, hits)."""
    out = np.empty((p.height, p.width, 4), np.float32)
    hits = np.empty((p.height, p.width), np.int32) if want_hits else None
    samples = c_uint64(0)
    check(lib.vr_render_to_host(ctx.h, volume.h, byref(p), out.ctypes.data,
                                hits.ctypes.data if want_hits else None,
                                ctypes.cast(byref(samples), c_void_p) if want_samples else None))
    return out, samples.value, hits


def render_to_host_static(ctx, volume, p, want_hits=False, want_samples=True):
    """render_to_host through vr_render_to_host_static: no device allocation, frame marched into the library's static
    scratch array (at most static_max_pixels() pixels; calls that overlap on one GPU race on it)."""
    out = np.empty((p.height, p.width, 4), np.float32)
    hits = np.empty((p.height, p.width), np.int32) if want_hits else None
    samples = c_uint64(0)
    check(lib.vr_render_to_host_static(ctx.h, volume.h, byref(p), out.ctypes.data,
                                       hits.ctypes.data if want_hits else None,
                                       ctypes.cast(byref(samples), c_void_p) if want_samples else None))
    return out, samples.value, hits


def static_max_pixels():
    return int(lib.vr_static_scratch_max_pixels())


def render_device(ctx, volume, p, out_ptr, hit_ptr=None, samples_ptr=None):
    check(lib.vr_render(ctx.h, volume.h, byref(p), c_void_p(out_ptr), c_void_p(hit_ptr) if hit_ptr else None,
                        c_void_p(samples_ptr) if samples_ptr else None))


def render_compressed_device(ctx, plan, p, out_ptr, hit_ptr=None, samples_ptr=None):
    check(lib.vr_render_compressed(ctx.h, plan.h, byref(p), c_void_p(out_ptr), c_void_p(hit_ptr) if hit_ptr else None,
                                   c_void_p(samples_ptr) if samples_ptr else None))


def render_compressed_lazy_device(ctx, plan, p, out_ptr, hit_ptr=None, samples_ptr=None):
    """Frame straight from the streams, decoding only the tiles the frame can read (vr_render_compressed_lazy).
    Returns {tiles, tiles_written, voxel_bytes, stream_bytes}."""
    st = (ctypes.c_int64 * 4)()
    check(lib.vr_render_compressed_lazy(ctx.h, plan.h, byref(p), c_void_p(out_ptr), c_void_p(hit_ptr) if hit_ptr else None,
                                        c_void_p(samples_ptr) if samples_ptr else None, st))
    return {"tiles": st[0], "tiles_written": st[1], "voxel_bytes": st[2], "stream_bytes": st[3]}


def lazy_tile_mask(streams, origins, dims, p):
    """Host-only: (needed uint8 [n, tiles], tile origins int32 [n, tiles, 3]) for a frame of `p` over a volume of `dims`
    (x, y, z) with brick b at origins[b] (vr_lazy_tile_mask)."""
    n = len(streams)
    hs = (c_void_p * n)(*[s.h for s in streams])
    org = np.ascontiguousarray(np.asarray(origins, np.int64).reshape(n, 3))
    dm = (ctypes.c_int64 * 3)(*[int(v) for v in dims])
    nt = c_int32(0)
    check(lib.vr_lazy_tile_mask(hs, n, org.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), dm, byref(p), None, None, byref(nt)))
    needed = np.zeros((n, nt.value), np.uint8)
    boxes = np.zeros((n, nt.value, 3), np.int32)
    check(lib.vr_lazy_tile_mask(hs, n, org.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), dm, byref(p), needed.ctypes.data,
                                boxes.ctypes.data, byref(nt)))
    return needed, boxes


def stream_tile_bounds(stream):
    """(mins, maxs, tile_dims): exact bounds of the decoded voxels of every decode tile, from the index alone."""
    td = (c_int32 * 3)()
    check(lib.vr_stream_tile_bounds(stream.h, None, None, 0, td))
    # number of tiles = brick voxels / tile voxels
    hdr = stream.header
    nt = int(hdr.X * hdr.Y * hdr.Z) // (td[0] * td[1] * td[2])
    mins = np.empty(nt, np.uint8)
    maxs = np.empty(nt, np.uint8)
    check(lib.vr_stream_tile_bounds(stream.h, mins.ctypes.data, maxs.ctypes.data, nt, td))
    return mins, maxs, tuple(td)


def smoke_check(vr, oa, decoded):
    """Used by __graft_entry__.smoke(): one 64x48 composite frame of the decoded brick vs the oracle."""
    ctx = vr.Context(0)
    Z, Y, X = decoded.shape
    vol = vr.Volume(ctx, X, Y, Z)
    vol.upload_brick(decoded, (0, 0, 0))
    p = default_params(64, 48, COMPOSITE, step=(1 / 64, 1 / 64, 1 / 32))
    got, S, _ = render_to_host(ctx, vol, p)
    want, S2, _ = oa.oracle_render(decoded, oracle_param_blob(p), threads=4)
    assert S == S2, (S, S2)
    assert np.abs(got - want).max() <= 1.0 / 255.0
