"""Host-side mirror of the reference codec interface on top of the C ABI.

`VolumeKdtree` keeps the reference's method names and argument meaning
(/root/reference/volume_renderer/VolumeKdtree_recover.h:103-175): build, levelCut, save, open,
setErrorTolerance, setMaxEpochs, measureMaxError, measureMeanError, queryError, and the public
fields tree / distanceMap / maxTreeDepth / origTreeDepth / numActiveNodes / X / Y / Z.
levelCut runs on the GPU (stage 1 of the hot path); build is the CPU reference algorithm.
"""
import ctypes
from ctypes import byref, c_int64, c_void_p

import numpy as np

from . import _lib
from ._lib import check, lib


class Context:
    """One device + CUDA stream (vr_context)."""

    def __init__(self, device=0, cuda_stream=None):
        self.h = c_void_p()
        check(lib.vr_context_create(int(device), byref(self.h)))
        self.device = int(device)
        if cuda_stream is not None:
            check(lib.vr_context_set_stream(self.h, c_void_p(int(cuda_stream))))

    def synchronize(self):
        check(lib.vr_context_synchronize(self.h))

    def wait_for(self, other):
        """Later work on this context waits (on the device) for everything already submitted to `other`."""
        check(lib.vr_context_wait_for(self.h, other.h))

    def download(self, dst_host_ptr, src_dev_ptr, nbytes):
        """Asynchronous device -> host copy on this context's stream."""
        check(lib.vr_context_download(self.h, c_void_p(int(dst_host_ptr)), c_void_p(int(src_dev_ptr)), int(nbytes)))

    def timer_begin(self, slot=0):
        check(lib.vr_timer_begin(self.h, slot))

    def timer_end(self, slot=0):
        check(lib.vr_timer_end(self.h, slot))

    def timer_ms(self, slot=0):
        ms = ctypes.c_float()
        check(lib.vr_timer_elapsed_ms(self.h, slot, byref(ms)))
        return ms.value

    def close(self):
        if self.h:
            lib.vr_context_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Stream:
    """One kd-tree stream (vr_stream): container fields + derived decode index."""

    def __init__(self, handle):
        self.h = handle

    @classmethod
    def from_memory(cls, header, distance_map, bits):
        dm = np.ascontiguousarray(distance_map, dtype=np.uint8)
        b = np.ascontiguousarray(bits, dtype=np.uint8)
        h = c_void_p()
        check(lib.vr_stream_from_memory(byref(header), dm.ctypes.data, b.ctypes.data, b.size, byref(h)))
        return cls(h)

    @classmethod
    def from_nibbles(cls, header, distance_map, nibbles):
        """MidRangeTree's 4-bit/node shader stream (convertToByteArray) -> (decodable midpoint Stream, range bits)."""
        dm = np.ascontiguousarray(distance_map, dtype=np.uint8)
        nb = np.ascontiguousarray(nibbles, dtype=np.uint8)
        rng = np.zeros((header.numActiveNodes + 3) // 4, np.uint8)
        h = c_void_p()
        check(lib.vr_stream_from_nibbles(byref(header), dm.ctypes.data, nb.ctypes.data, nb.size, byref(h), rng.ctypes.data))
        return cls(h), rng

    @classmethod
    def open(cls, path):
        h = c_void_p()
        check(lib.vr_stream_open(str(path).encode(), byref(h)))
        return cls(h)

    @classmethod
    def compress(cls, brick, tolerance=6, max_epochs=5):
        """brick: uint8 array shaped (Z, Y, X) (x fastest).  Returns (Stream, BuildStats)."""
        a = np.ascontiguousarray(brick, dtype=np.uint8)
        assert a.ndim == 3
        Z, Y, X = a.shape
        h = c_void_p()
        st = _lib.BuildStats()
        check(lib.vr_kd_compress(a.ctypes.data, X, Y, Z, int(tolerance), int(max_epochs), byref(h), byref(st)))
        return cls(h), st

    @classmethod
    def compress_device(cls, ctx, brick, tolerance=6, max_epochs=5):
        """Same stream as compress(), produced on the GPU (csrc/kd_build.cu); brick is a host array (Z, Y, X)."""
        a = np.ascontiguousarray(brick, dtype=np.uint8)
        assert a.ndim == 3
        Z, Y, X = a.shape
        h = c_void_p()
        st = _lib.BuildStats()
        check(lib.vr_kd_compress_device(ctx.h, a.ctypes.data, X, Y, Z, int(tolerance), int(max_epochs), byref(h), byref(st)))
        return cls(h), st

    @classmethod
    def compress_volume(cls, ctx, volume, origin, shape, tolerance=6, max_epochs=5):
        """GPU producer on the (X, Y, Z) box at voxel `origin` (x, y, z) of a device Volume.
        Returns (Stream, BuildStats, (scan passes of the running-mean fixed point: total, worst level))."""
        import ctypes
        org = (ctypes.c_int64 * 3)(*[int(v) for v in origin])
        it = (ctypes.c_int64 * 2)(0, 0)
        h = c_void_p()
        st = _lib.BuildStats()
        X, Y, Z = [int(v) for v in shape]
        check(lib.vr_kd_compress_volume(ctx.h, volume.h, org, X, Y, Z, int(tolerance), int(max_epochs), byref(h), byref(st), it))
        return cls(h), st, (int(it[0]), int(it[1]))

    def save(self, path):
        check(lib.vr_stream_save(self.h, str(path).encode()))

    # ---- indexed ("random-access") container, csrc/kd_kdx.cpp ----
    def save_indexed(self, path):
        check(lib.vr_stream_save_indexed(self.h, str(path).encode()))

    def serialize_indexed(self):
        """The bytes save_indexed() writes, as a uint8 array."""
        n = c_int64()
        check(lib.vr_stream_serialize_indexed(self.h, None, 0, byref(n)))
        out = np.empty(n.value, np.uint8)
        check(lib.vr_stream_serialize_indexed(self.h, out.ctypes.data, out.size, byref(n)))
        return out

    @classmethod
    def open_indexed(cls, path, tile_mask=None, verify=False):
        """Opens an indexed container without walking the stream.  tile_mask (uint8[tiles]): fetch only those tiles' symbols.
        Returns the Stream; .bytes_read = bytes fetched from the file."""
        m = None if tile_mask is None else np.ascontiguousarray(tile_mask, dtype=np.uint8)
        h = c_void_p()
        nread = c_int64()
        check(lib.vr_stream_open_indexed(str(path).encode(), None if m is None else m.ctypes.data, int(bool(verify)), byref(h), byref(nread)))
        s = cls(h)
        s.bytes_read = nread.value
        return s

    @classmethod
    def from_memory_indexed(cls, image, tile_mask=None, verify=False):
        a = np.ascontiguousarray(image, dtype=np.uint8)
        m = None if tile_mask is None else np.ascontiguousarray(tile_mask, dtype=np.uint8)
        h = c_void_p()
        check(lib.vr_stream_from_memory_indexed(a.ctypes.data, a.size, None if m is None else m.ctypes.data, int(bool(verify)), byref(h)))
        return cls(h)

    def index_info(self):
        out = (c_int64 * 6)()
        check(lib.vr_stream_index_info(self.h, out))
        return {"tiles": out[0], "live_tiles": out[1], "leaves_per_tile": out[2], "subset": bool(out[3]), "index_bytes": out[4],
                "max_symbols_per_leaf": out[5]}

    _SECTIONS = {"tileLeafBlock": (1, np.uint32), "tileVal": (2, np.uint8), "tileMin": (3, np.uint8), "tileMax": (4, np.uint8),
                 "blockSymBegin": (5, np.uint32), "blockSymEnd": (6, np.uint32), "leafOff": (7, np.uint32), "leafVal": (8, np.uint8)}

    def index_section(self, name):
        """Copy of one array of the decode index (names as in csrc/kd_stream.h KdIndex)."""
        which, dt = self._SECTIONS[name]
        p = c_void_p()
        n = c_int64()
        check(lib.vr_stream_index_section(self.h, which, byref(p), byref(n)))
        if n.value == 0:
            return np.zeros(0, dt)
        raw = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), shape=(n.value,)).copy()
        return raw.view(dt)

    def index_arrays(self):
        return {k: self.index_section(k) for k in self._SECTIONS}

    @property
    def header(self):
        hd = _lib.StreamHeader()
        check(lib.vr_stream_get_header(self.h, byref(hd)))
        return hd

    @property
    def distance_map(self):
        out = np.zeros(self.header.maxTreeDepth + 1, np.uint8)
        check(lib.vr_stream_get_distance_map(self.h, out.ctypes.data, out.size))
        return out

    @property
    def bits(self):
        p = c_void_p()
        n = c_int64()
        check(lib.vr_stream_get_bits(self.h, byref(p), byref(n)))
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), shape=(n.value,)).copy()

    def validate(self):
        n = c_int64()
        check(lib.vr_stream_validate(self.h, byref(n)))
        return n.value

    def prepare(self):
        check(lib.vr_stream_prepare(self.h))
        return self

    def pin(self):
        check(lib.vr_stream_pin(self.h))
        return self

    def device_bytes(self):
        out = (c_int64 * 2)()
        check(lib.vr_stream_device_bytes(self.h, out))
        return out[0], out[1]

    def upload(self, ctx):
        check(lib.vr_stream_upload(ctx.h, self.h))
        return self

    def close(self):
        if self.h:
            lib.vr_stream_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_streams(paths, ctx=None, threads=0):
    """VolumeKdtree::open for many files at once (vr_streams_load): worker threads open + index the containers while the
    calling thread pins and uploads them to ctx's device in order.  ctx=None: open + prepare only."""
    n = len(paths)
    arr = (ctypes.c_char_p * n)(*[str(p).encode() for p in paths])
    out = (c_void_p * n)()
    check(lib.vr_streams_load(ctx.h if ctx is not None else None, arr, n, int(threads), out))
    return [Stream(c_void_p(out[i])) for i in range(n)]


class StreamArena:
    """vr_stream_arena: the images of a timestep's streams in one pinned buffer + one device buffer; upload() is ONE copy.
    Create it before the DecodePlans of these streams."""

    def __init__(self, ctx, streams):
        self.streams = list(streams)
        n = len(self.streams)
        self.h = c_void_p()
        check(lib.vr_stream_arena_create(ctx.h, (c_void_p * n)(*[s.h for s in self.streams]), n, byref(self.h)))

    def upload(self, ctx):
        check(lib.vr_stream_arena_upload(ctx.h, self.h))

    @property
    def nbytes(self):
        b = c_int64()
        check(lib.vr_stream_arena_bytes(self.h, byref(b)))
        return b.value

    def close(self):
        if self.h:
            lib.vr_stream_arena_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def mrt_save(path, header, distance_map, distance_map_range, tree_bits, range_bits):
    """MidRangeTree::save (MidRangeTree.cpp:753-786): byte-identical file from the class's public fields."""
    dm = np.ascontiguousarray(distance_map, dtype=np.uint8)
    dmr = np.ascontiguousarray(distance_map_range, dtype=np.uint8)
    tb = np.ascontiguousarray(tree_bits, dtype=np.uint8)
    rb = np.ascontiguousarray(range_bits, dtype=np.uint8)
    assert tb.size == rb.size and dm.size == dmr.size == header.maxTreeDepth + 1
    check(lib.vr_mrt_save(str(path).encode(), byref(header), dm.ctypes.data, dmr.ctypes.data, tb.ctypes.data, rb.ctypes.data, tb.size))


def mrt_open(path):
    """MidRangeTree::open without its 4-byte misalignment: (midpoint Stream, distanceMap_range, tree_range bits)."""
    h = c_void_p()
    n = c_int64()
    check(lib.vr_mrt_open(str(path).encode(), byref(h), None, 0, None, 0, byref(n)))
    lib.vr_stream_destroy(h)
    dmr = np.zeros(64, np.uint8)
    rb = np.zeros(n.value, np.uint8)
    h = c_void_p()
    check(lib.vr_mrt_open(str(path).encode(), byref(h), dmr.ctypes.data, dmr.size, rb.ctypes.data, rb.size, byref(n)))
    s = Stream(h)
    return s, dmr[: s.header.maxTreeDepth + 1].copy(), rb


def nibbles_pack(tree_bits, range_bits, n):
    """MidRangeTree::convertToByteArray (MidRangeTree.cpp:1095-1128): the 4-bit/node stream of n nodes."""
    t = np.ascontiguousarray(tree_bits, np.uint8); r = np.ascontiguousarray(range_bits, np.uint8)
    L = c_int64()
    check(lib.vr_nibbles_pack(t.ctypes.data, r.ctypes.data, int(n), None, 0, byref(L)))
    out = np.zeros(L.value, np.uint8)
    check(lib.vr_nibbles_pack(t.ctypes.data, r.ctypes.data, int(n), out.ctypes.data, out.size, byref(L)))
    return out


def nibbles_unpack(nibbles, n):
    nb = np.ascontiguousarray(nibbles, np.uint8)
    t = np.zeros((n + 3) // 4, np.uint8); r = np.zeros((n + 3) // 4, np.uint8)
    check(lib.vr_nibbles_unpack(nb.ctypes.data, nb.size, int(n), t.ctypes.data, r.ctypes.data))
    return t, r


class Volume:
    """Device slab (vr_volume), x fastest; numpy view order is (Z, Y, X)."""

    def __init__(self, ctx, SX, SY, SZ, device_ptr=None):
        self.ctx = ctx
        self.h = c_void_p()
        self.dims = (int(SX), int(SY), int(SZ))
        if device_ptr is None:
            check(lib.vr_volume_create(ctx.h, SX, SY, SZ, byref(self.h)))
        else:
            check(lib.vr_volume_wrap(ctx.h, c_void_p(int(device_ptr)), SX, SY, SZ, byref(self.h)))

    @property
    def device_ptr(self):
        p = c_void_p()
        check(lib.vr_volume_device_ptr(self.h, byref(p)))
        return p.value

    def download(self):
        SX, SY, SZ = self.dims
        out = np.empty((SZ, SY, SX), np.uint8)
        check(lib.vr_volume_download(self.ctx.h, self.h, out.ctypes.data))
        return out

    def set_cell_origin(self, origin):
        """Global voxel coordinate of this slab's voxel (0,0,0): occupancy cells align to the global grid."""
        check(lib.vr_volume_set_cell_origin(self.h, (c_int64 * 3)(*[int(o) for o in origin])))
        return self

    def set_skipping(self, enable=True):
        """Empty-space skipping in the march (default on); frames are identical either way."""
        check(lib.vr_volume_set_skipping(self.h, 1 if enable else 0))
        return self

    def invalidate(self):
        check(lib.vr_volume_invalidate(self.h))

    def occupancy(self, which=0):
        """Per-cell maxima as numpy (nz, ny, nx): 0 raw (cells of 32x32x16), 1 + l = level l of the octree of
        dilated maxima (level 0 = dilated cells; each level halves the grid)."""
        dims = (ctypes.c_int32 * 3)()
        check(lib.vr_volume_occupancy_download(self.ctx.h, self.h, int(which), None, 0, dims))
        out = np.empty((dims[2], dims[1], dims[0]), np.uint8)
        check(lib.vr_volume_occupancy_download(self.ctx.h, self.h, int(which), out.ctypes.data, out.size, dims))
        return out

    def load_bricks(self, numBricks, I, J, K, timestep, brick_dims, find_source_file, brick_map=None, want_host=False):
        """VolumeReader::LoadBricksToTexture (VolumeReader.h:151-223): brick b's raw file find_source_file(b, timestep)
        goes to grid position brick_map[b] (default: fillVolumeBrickMap's b -> (b % I, (b // I) % J, b // (I*J))),
        file reads and H2D copies pipelined (vr_volume_load_bricks).  Returns (bricks, bytes[, host array (Z, Y, X)])."""
        keep = []

        def _find(b, t, _user):
            keep.append(ctypes.create_string_buffer(str(find_source_file(b, t)).encode()))
            return ctypes.addressof(keep[-1])

        cb = _lib.FIND_SOURCE_FILE(_find)
        bd = (c_int64 * 3)(*[int(v) for v in brick_dims])
        ijk = None
        if brick_map is not None:
            ijk = (c_int64 * (3 * numBricks))(*[int(v) for b in range(numBricks) for v in brick_map[b]])
        host = np.zeros((K * brick_dims[2], J * brick_dims[1], I * brick_dims[0]), np.uint8) if want_host else None
        st = (c_int64 * 2)()
        check(lib.vr_volume_load_bricks(self.ctx.h, self.h, int(numBricks), int(I), int(J), int(K), int(timestep), bd,
                                        ctypes.cast(cb, c_void_p), None, ijk, host.ctypes.data if want_host else None, st))
        return (st[0], st[1], host) if want_host else (st[0], st[1])

    def upload_brick(self, brick, origin):
        a = np.ascontiguousarray(brick, dtype=np.uint8)
        BZ, BY, BX = a.shape
        check(lib.vr_volume_upload_brick(self.ctx.h, self.h, a.ctypes.data, BX, BY, BZ, *[int(o) for o in origin]))
        self.ctx.synchronize()  # `a` may be a temporary

    def close(self):
        if self.h:
            lib.vr_volume_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DecodePlan:
    """Batched stage-1 decode of uploaded streams into a Volume (one kernel launch per run)."""

    def __init__(self, ctx, streams, origins, volume):
        self.ctx, self.streams, self.volume = ctx, list(streams), volume
        n = len(self.streams)
        arr = (c_void_p * n)(*[s.h for s in self.streams])
        org = (c_int64 * (3 * n))(*[int(v) for o in origins for v in o])
        self.h = c_void_p()
        check(lib.vr_decode_plan_create(ctx.h, arr, n, org, volume.h, byref(self.h)))

    def run(self):
        check(lib.vr_decode_plan_run(self.h))

    def run_tiles(self, needed):
        """Random-access decode: only the tiles with needed[brick, tile] != 0 (vr_decode_plan_run_tiles).  Returns stats."""
        m = np.ascontiguousarray(needed, dtype=np.uint8)
        st = (c_int64 * 4)()
        check(lib.vr_decode_plan_run_tiles(self.h, m.ctypes.data, st))
        return {"tiles": st[0], "tiles_written": st[1], "voxel_bytes": st[2], "stream_bytes": st[3]}

    def bytes(self):
        out = (c_int64 * 3)()
        check(lib.vr_decode_plan_bytes(self.h, out))
        return {"stream": out[0], "voxels": out[1], "index": out[2]}

    def kernel(self):
        """'persistent' (warp-specialised kernel, kd_decode.cu) or 'classic' (one CTA per tile)."""
        from ctypes import c_int32
        w = c_int32(0)
        check(lib.vr_decode_plan_kernel(self.h, byref(w)))
        return "persistent" if w.value == 0 else "classic"

    def close(self):
        if self.h:
            lib.vr_decode_plan_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class VolumeKdtree:
    """Reference-shaped codec object (VolumeKdtree_recover.h:49-176)."""

    def __init__(self, inData=None, x=0, y=0, z=0, ctx=None):
        self._data = None if inData is None else np.ascontiguousarray(inData, dtype=np.uint8).reshape(z, y, x)
        self.X, self.Y, self.Z = int(x), int(y), int(z)
        self.tolerance, self.maxEpochs = 6, 5  # defaults VolumeKdtree_recover.h:92-93,110-111
        self._stream = None
        self._ctx = ctx
        self._output = None
        self.stats = None

    # -- knobs (VolumeKdTree_recover.cpp:9-15)
    def setErrorTolerance(self, errorTolerance):
        self.tolerance = int(errorTolerance)

    def setMaxEpochs(self, epochs):
        self.maxEpochs = int(epochs)

    # -- build (VolumeKdTree_recover.cpp:17); the input array is left intact (SURVEY C-11)
    def build(self, useThreads=True, device=False):
        """VolumeKdtree::build (VolumeKdtree_recover.h:126).  device=True runs the GPU producer (csrc/kd_build.cu) on
        this object's context; both give the byte-identical stream of the reference."""
        if self._data is None:
            raise _lib.VrError(_lib.VR_ERR_STATE, "build: no input volume")
        if device:
            if self._ctx is None:
                raise _lib.VrError(_lib.VR_ERR_STATE, "build(device=True) needs a Context")
            self._stream, self.stats = Stream.compress_device(self._ctx, self._data, self.tolerance, self.maxEpochs)
        else:
            self._stream, self.stats = Stream.compress(self._data, self.tolerance, self.maxEpochs)
        return self

    def _need(self):
        if self._stream is None:
            raise _lib.VrError(_lib.VR_ERR_STATE, "no tree: call build() or open() first")
        return self._stream

    # -- public fields of the reference class
    @property
    def stream(self):
        return self._need()

    @property
    def tree(self):
        return self._need().bits

    @property
    def distanceMap(self):
        return self._need().distance_map

    @property
    def maxTreeDepth(self):
        return self._need().header.maxTreeDepth

    @property
    def origTreeDepth(self):
        return self._need().header.origTreeDepth

    @property
    def numActiveNodes(self):
        return self._need().header.numActiveNodes

    def save(self, filename):
        self._need().save(filename)

    def open(self, filename):
        self._stream = Stream.open(filename)
        hd = self._stream.header
        self.X, self.Y, self.Z = hd.X, hd.Y, hd.Z
        return self

    # -- stage 1 on the GPU (VolumeKdTree_recover.cpp:726)
    def levelCut(self, cutDepth, outData=None):
        s = self._need()
        if self._ctx is None:
            self._ctx = Context(0)
        out = np.empty((self.Z, self.Y, self.X), np.uint8) if outData is None else outData
        check(lib.vr_kd_levelcut(self._ctx.h, s.h, int(cutDepth), out.ctypes.data))
        self._output = out
        return out

    # -- error queries (VolumeKdTree_recover.cpp:386-411); usable here because build keeps the input
    def measureMaxError(self):
        return int(np.abs(self._output.astype(np.int16) - self._data.astype(np.int16)).max())

    def measureMeanError(self):
        return float(np.abs(self._output.astype(np.int16) - self._data.astype(np.int16)).mean())

    def queryError(self):
        return np.abs(self._output.astype(np.int16) - self._data.astype(np.int16)).astype(np.uint8)
