"""ctypes binding of libvr_b200.so (the C ABI declared in include/vr_b200.h).

The library is built in-tree by build.sh / __graft_entry__.build().  There is no fallback: if the
shared object is missing, importing this module raises.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_uint8, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvr_b200.so")

VR_OK, VR_ERR_ARG, VR_ERR_IO, VR_ERR_FORMAT, VR_ERR_CUDA, VR_ERR_UNSUPPORTED, VR_ERR_STATE = range(7)


class VrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vr_b200 error {code}: {msg}")
        self.code = code


class StreamHeader(ctypes.Structure):
    """88-byte container header (VolumeKdTree_recover.cpp:535-542)."""
    _pack_ = 1
    _fields_ = [("rootMin", c_int64 * 3), ("rootMax", c_int64 * 3), ("maxTreeDepth", c_int32),
                ("origTreeDepth", c_int32), ("X", c_int64), ("Y", c_int64), ("Z", c_int64),
                ("numActiveNodes", c_int64)]


class BuildStats(ctypes.Structure):
    _fields_ = [("maxErrorBeforeGrowth", c_int32), ("maxError", c_int32), ("meanL1", c_double),
                ("meanL2", c_double), ("zeroRunRewrites", c_int64)]


if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: run ./build.sh (or __graft_entry__.build()) first; "
                      "there is no CPU fallback for the CUDA path")

lib = ctypes.CDLL(LIB_PATH)

_SIGS = {
    "vr_last_error": (c_char_p, []),
    "vr_version": (c_char_p, []),
    "vr_context_create": (c_int, [c_int, POINTER(c_void_p)]),
    "vr_context_destroy": (c_int, [c_void_p]),
    "vr_context_set_stream": (c_int, [c_void_p, c_void_p]),
    "vr_context_synchronize": (c_int, [c_void_p]),
    "vr_context_wait_for": (c_int, [c_void_p, c_void_p]),
    "vr_context_download": (c_int, [c_void_p, c_void_p, c_void_p, c_int64]),
    "vr_timer_begin": (c_int, [c_void_p, c_int]),
    "vr_timer_end": (c_int, [c_void_p, c_int]),
    "vr_timer_elapsed_ms": (c_int, [c_void_p, c_int, POINTER(ctypes.c_float)]),
    "vr_stream_from_memory": (c_int, [POINTER(StreamHeader), c_void_p, c_void_p, c_int64, POINTER(c_void_p)]),
    "vr_stream_open": (c_int, [c_char_p, POINTER(c_void_p)]),
    "vr_stream_save": (c_int, [c_void_p, c_char_p]),
    "vr_stream_destroy": (c_int, [c_void_p]),
    "vr_stream_save_indexed": (c_int, [c_void_p, c_char_p]),
    "vr_stream_open_indexed": (c_int, [c_char_p, c_void_p, c_int32, POINTER(c_void_p), POINTER(c_int64)]),
    "vr_stream_from_memory_indexed": (c_int, [c_void_p, c_int64, c_void_p, c_int32, POINTER(c_void_p)]),
    "vr_stream_serialize_indexed": (c_int, [c_void_p, c_void_p, c_int64, POINTER(c_int64)]),
    "vr_stream_index_info": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_stream_index_section": (c_int, [c_void_p, c_int32, POINTER(c_void_p), POINTER(c_int64)]),
    "vr_stream_get_header": (c_int, [c_void_p, POINTER(StreamHeader)]),
    "vr_stream_get_distance_map": (c_int, [c_void_p, c_void_p, c_int32]),
    "vr_stream_get_bits": (c_int, [c_void_p, POINTER(c_void_p), POINTER(c_int64)]),
    "vr_stream_validate": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_stream_prepare": (c_int, [c_void_p]),
    "vr_stream_device_bytes": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_stream_pin": (c_int, [c_void_p]),
    "vr_stream_upload": (c_int, [c_void_p, c_void_p]),
    "vr_nibbles_pack": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, POINTER(c_int64)]),
    "vr_nibbles_unpack": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "vr_stream_from_nibbles": (c_int, [POINTER(StreamHeader), c_void_p, c_void_p, c_int64, POINTER(c_void_p), c_void_p]),
    "vr_mrt_save": (c_int, [c_char_p, POINTER(StreamHeader), c_void_p, c_void_p, c_void_p, c_void_p, c_int64]),
    "vr_mrt_open": (c_int, [c_char_p, POINTER(c_void_p), c_void_p, c_int32, c_void_p, c_int64, POINTER(c_int64)]),
    "vr_kd_compress": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, POINTER(c_void_p), POINTER(BuildStats)]),
    "vr_peer_create": (c_int, [c_void_p, c_int32, c_int32, c_int64, POINTER(c_void_p)]),
    "vr_peer_destroy": (c_int, [c_void_p]),
    "vr_peer_handle": (c_int, [c_void_p, c_void_p]),
    "vr_peer_open": (c_int, [c_void_p, c_void_p]),
    "vr_peer_raw": (c_int, [c_void_p, POINTER(c_void_p)]),
    "vr_peer_open_raw": (c_int, [c_void_p, POINTER(c_void_p)]),
    "vr_peer_window": (c_int, [c_void_p, c_int32, POINTER(c_void_p)]),
    "vr_peer_put": (c_int, [c_void_p, c_int32, c_int32, POINTER(c_int32), POINTER(c_int64), POINTER(c_void_p), POINTER(c_int64)]),
    "vr_peer_wait": (c_int, [c_void_p, c_int32, c_int32, POINTER(c_int32)]),
    "vr_peer_reduce": (c_int, [c_void_p, c_int32, c_int32, POINTER(c_void_p), c_void_p, c_int64]),
    "vr_peer_error": (c_int, [c_void_p, POINTER(c_int32)]),
    "vr_peer_partial_composite": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "vr_peer_resolve_composite": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p]),
    "vr_peer_finish_composite": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_int64, c_int64]),
    "vr_peer_finish_composite_to": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_int64, c_int64, c_int32]),
    "vr_kd_compress_device": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, POINTER(c_void_p), POINTER(BuildStats)]),
    "vr_kd_compress_volume": (c_int, [c_void_p, c_void_p, POINTER(c_int64), c_int64, c_int64, c_int64, c_int, c_int, POINTER(c_void_p),
                                      POINTER(BuildStats), POINTER(c_int64)]),
    "vr_volume_create": (c_int, [c_void_p, c_int64, c_int64, c_int64, POINTER(c_void_p)]),
    "vr_volume_wrap": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, POINTER(c_void_p)]),
    "vr_volume_destroy": (c_int, [c_void_p]),
    "vr_volume_device_ptr": (c_int, [c_void_p, POINTER(c_void_p)]),
    "vr_volume_dims": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_volume_download": (c_int, [c_void_p, c_void_p, c_void_p]),
    "vr_volume_upload_brick": (c_int, [c_void_p, c_void_p, c_void_p] + [c_int64] * 6),
    "vr_volume_set_cell_origin": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_volume_set_skipping": (c_int, [c_void_p, c_int]),
    "vr_volume_invalidate": (c_int, [c_void_p]),
    "vr_volume_occupancy_download": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_int64, POINTER(c_int32)]),
    "vr_decode_plan_create": (c_int, [c_void_p, POINTER(c_void_p), c_int32, POINTER(c_int64), c_void_p, POINTER(c_void_p)]),
    "vr_decode_plan_run": (c_int, [c_void_p]),
    "vr_decode_plan_run_tiles": (c_int, [c_void_p, c_void_p, POINTER(c_int64)]),
    "vr_decode_plan_destroy": (c_int, [c_void_p]),
    "vr_decode_plan_bytes": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_decode_plan_kernel": (c_int, [c_void_p, POINTER(c_int32)]),
    "vr_debug_set": (c_int, [c_char_p, c_int]),
    "vr_debug_get": (c_int, [c_char_p, POINTER(c_int)]),
    "vr_kd_levelcut": (c_int, [c_void_p, c_void_p, c_int32, c_void_p]),
    "vr_volume_load_bricks": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int32, POINTER(c_int64), c_void_p, c_void_p,
                                      POINTER(c_int64), c_void_p, POINTER(c_int64)]),
    "vr_stream_arena_create": (c_int, [c_void_p, POINTER(c_void_p), c_int32, POINTER(c_void_p)]),
    "vr_stream_arena_upload": (c_int, [c_void_p, c_void_p]),
    "vr_stream_arena_bytes": (c_int, [c_void_p, POINTER(c_int64)]),
    "vr_stream_arena_destroy": (c_int, [c_void_p]),
    "vr_streams_load": (c_int, [c_void_p, POINTER(c_char_p), c_int32, c_int32, POINTER(c_void_p)]),
}
FIND_SOURCE_FILE = ctypes.CFUNCTYPE(c_void_p, c_int32, c_int32, c_void_p)  # returns the address of a NUL-terminated path

EXPORTED = sorted(_SIGS)
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)  # AttributeError here = header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args


def check(rc):
    if rc != VR_OK:
        raise VrError(rc, lib.vr_last_error().decode("utf-8", "replace"))
    return rc
