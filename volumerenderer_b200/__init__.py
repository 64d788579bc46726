"""volumerenderer_b200 -- B200-native kd-tree stream decode + ray-marching (hot path of
AugmentariumLab/VolumeRenderer) behind a C ABI (include/vr_b200.h, libvr_b200.so).

Importing this package loads the CUDA library; it raises if the library was not built.
"""
from . import _lib, render, sortlast, synth  # noqa: F401
from ._lib import VrError  # noqa: F401
from .kdtree import Context, DecodePlan, Stream, Volume, VolumeKdtree, load_streams, nibbles_pack, nibbles_unpack  # noqa: F401
from .loader import VolumeReader  # noqa: F401
