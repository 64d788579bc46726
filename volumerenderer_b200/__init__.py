"""volumerenderer_b200 -- B200-native kd-tree stream decode + ray-marching (hot path of
AugmentariumLab/VolumeRenderer) behind a C ABI (include/vr_b200.h, libvr_b200.so).

Importing this package loads the CUDA library; it raises if the library was not built.
"""
from . import _lib, render, sortlast, synth  # noqa: F401
from ._lib import VrError  # noqa: F401
from .kdtree import Context, DecodePlan, Stream, StreamArena, Volume, VolumeKdtree, load_streams, mrt_open, mrt_save, nibbles_pack, nibbles_unpack  # noqa: F401
from .loader import VolumeReader  # noqa: F401


def debug_set(key, value):
    """vr_debug_set: explicit A/B knobs of the decode launch ("decode_kernel", "decode_only", "fill_attach")."""
    _lib.check(_lib.lib.vr_debug_set(key.encode(), int(value)))
