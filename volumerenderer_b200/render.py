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
for _n in ("vr_render", "vr_render_to_host", "vr_render_compressed"):
    getattr(_lib.lib, _n).restype = ctypes.c_int
    getattr(_lib.lib, _n).argtypes = [c_void_p, c_void_p, ctypes.POINTER(RenderParams), c_void_p, c_void_p, c_void_p]
_lib.EXPORTED[:] = sorted(set(_lib.EXPORTED) | {"vr_render_defaults", "vr_camera_basis", "vr_render",
                                                "vr_render_to_host", "vr_render_compressed"})


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


def render_device(ctx, volume, p, out_ptr, hit_ptr=None, samples_ptr=None):
    check(lib.vr_render(ctx.h, volume.h, byref(p), c_void_p(out_ptr), c_void_p(hit_ptr) if hit_ptr else None,
                        c_void_p(samples_ptr) if samples_ptr else None))


def render_compressed_device(ctx, plan, p, out_ptr, hit_ptr=None, samples_ptr=None):
    check(lib.vr_render_compressed(ctx.h, plan.h, byref(p), c_void_p(out_ptr), c_void_p(hit_ptr) if hit_ptr else None,
                                   c_void_p(samples_ptr) if samples_ptr else None))


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
