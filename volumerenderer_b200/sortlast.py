"""Sort-last rendering of a volume sharded over GPUs: host-side logic above the C ABI.

The reference renders one 3-D texture on one GPU (main.cpp:290,404).  Here every rank (one process per
GPU) holds a box of that texture -- its own bricks decoded in place plus a halo -- and a frame is

  composite  (raycaster.frag):  pass 1 per-slab segment partials (C_seg, T_seg)  ->  all-gather of T_seg
                                ->  pass 2 per-rank (C, A) term  ->  all-reduce(sum)  ->  colour map
  isosurface (isosurface.frag): pass 1 first hit among the slab's samples  ->  all-reduce(min) of the hit
                                index  ->  mask  ->  all-reduce(sum) of the colours

The kernels are the C ABI's vr_partial_* / vr_resolve_* / vr_finish_* (csrc/render_sortlast.cu); the
exchange is torch.distributed (NCCL on GPUs; the same code runs over gloo with CPU tensors in the tests,
with a CPU stand-in for the kernels supplied by the test).  Nothing here touches the oracle.
"""
import ctypes
from ctypes import byref, c_int32, c_int64, c_void_p

import numpy as np

from . import _lib
from ._lib import check, lib
from .render import COMPOSITE, RenderParams

HIT_NONE = 0x7FFFFFFF


class SlabDesc(ctypes.Structure):
    """vr_slab_desc: where a rank's device slab sits in the whole texture."""
    _fields_ = [("global_", c_int64 * 3), ("origin", c_int64 * 3), ("own_lo", c_int64 * 3), ("own_hi", c_int64 * 3)]

    def stored(self, halo):
        """(origin, extent) of the device slab for a halo of `halo` voxels per axis and side."""
        o = [max(0, self.own_lo[a] - halo[a]) for a in range(3)]
        e = [min(self.global_[a], self.own_hi[a] + halo[a]) - o[a] for a in range(3)]
        return o, e


_P = ctypes.POINTER
_SIGS = {
    "vr_sortlast_create": [c_void_p, _P(SlabDesc), c_int32, c_int32, _P(c_void_p)],
    "vr_sortlast_destroy": [c_void_p],
    "vr_sortlast_order": [c_void_p, c_int32, _P(c_int32)],
    "vr_sortlast_halo": [_P(RenderParams), _P(c_int64), _P(c_int64)],
    "vr_partial_composite": [c_void_p, c_void_p, c_void_p, _P(RenderParams), c_void_p, c_void_p, c_void_p],
    "vr_partial_iso": [c_void_p, c_void_p, c_void_p, _P(RenderParams), c_void_p, c_void_p, c_void_p],
    "vr_resolve_composite": [c_void_p, c_void_p, c_void_p, _P(RenderParams), c_void_p, c_void_p, c_void_p, c_void_p],
    "vr_finish_composite": [c_void_p, _P(RenderParams), c_void_p, c_void_p],
    "vr_resolve_iso": [c_void_p, c_void_p, _P(RenderParams), c_void_p, c_void_p, c_void_p],
    "vr_volume_pack_box": [c_void_p, c_void_p, _P(c_int64), _P(c_int64), c_void_p],
    "vr_volume_unpack_box": [c_void_p, c_void_p, _P(c_int64), _P(c_int64), c_void_p],
    "vr_peer_tin_composite": [c_void_p, c_void_p, c_int32, c_void_p, _P(RenderParams), c_int64, c_int64],
    "vr_resolve_composite_tin": [c_void_p, c_void_p, c_void_p, _P(RenderParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "vr_sortlast_frame_window_bytes": [_P(SlabDesc), c_int32, c_int32, c_int32, _P(c_int64), c_int32, _P(c_int64)],
    "vr_sortlast_frame_create": [c_void_p, c_void_p, c_void_p, c_int32, c_int32, _P(c_int64), c_int32, _P(c_void_p)],
    "vr_sortlast_frame_destroy": [c_void_p],
    "vr_sortlast_exchange_halos": [c_void_p, c_void_p, c_int32, c_int32],
    "vr_sortlast_render": [c_void_p, c_void_p, _P(RenderParams), _P(c_void_p), c_void_p],
}
for _n, _a in _SIGS.items():
    getattr(lib, _n).restype = ctypes.c_int
    getattr(lib, _n).argtypes = _a
_lib.EXPORTED[:] = sorted(set(_lib.EXPORTED) | set(_SIGS))


# ---------------------------------------------------------------------------------------------------
# layout
def required_halo(params, global_dims):
    """Voxels per axis and side a slab must hold around its owned box (vr_sortlast_halo)."""
    g = (c_int64 * 3)(*[int(v) for v in global_dims])
    h = (c_int64 * 3)()
    check(lib.vr_sortlast_halo(byref(params), g, h))
    return [int(v) for v in h]


def slab_layout(global_dims, n_ranks, axis=1, halo=(1, 1, 1)):
    """Partition the texture into n_ranks slabs along `axis` (y-slabs by brick row).
    Returns (descs, stored) with stored[r] = (origin, extent) of rank r's device slab."""
    G = [int(v) for v in global_dims]
    if G[axis] % n_ranks:
        raise ValueError(f"extent {G[axis]} along axis {axis} does not divide into {n_ranks} slabs")
    w = G[axis] // n_ranks
    descs, stored = [], []
    for r in range(n_ranks):
        d = SlabDesc()
        lo, hi = [0, 0, 0], list(G)
        lo[axis], hi[axis] = r * w, (r + 1) * w
        d.global_[:] = G
        d.own_lo[:] = lo
        d.own_hi[:] = hi
        o, e = d.stored(halo)
        d.origin[:] = o
        descs.append(d)
        stored.append((o, e))
    return descs, stored


def halo_transfers(descs, stored):
    """Boxes that must travel between ranks so every device slab is complete: for each ordered pair
    (src, dst) the part of dst's stored box that src owns.  Returns [(src, dst, global_lo, extent)]."""
    out = []
    for dst, (o, e) in enumerate(stored):
        for src, d in enumerate(descs):
            if src == dst:
                continue
            lo = [max(o[a], d.own_lo[a]) for a in range(3)]
            hi = [min(o[a] + e[a], d.own_hi[a]) for a in range(3)]
            if all(hi[a] > lo[a] for a in range(3)):
                out.append((src, dst, lo, [hi[a] - lo[a] for a in range(3)]))
    return out


def halo_staging_bytes(descs, stored):
    """Bytes of halo staging a PeerExchange window needs (the largest sum of incoming boxes over the ranks)."""
    need = {}
    for src, dst, lo, ext in halo_transfers(descs, stored):
        need[dst] = need.get(dst, 0) + (int(ext[0]) * int(ext[1]) * int(ext[2]) + 15) // 16 * 16
    return max(need.values()) if need else 0


class SortLast:
    """vr_sortlast: the slab geometry of all ranks + the visibility orders, bound to one rank."""

    def __init__(self, ctx, descs, rank):
        self.ctx, self.n, self.rank = ctx, len(descs), int(rank)
        arr = (SlabDesc * self.n)(*descs)
        self.h = c_void_p()
        check(lib.vr_sortlast_create(ctx.h if ctx is not None else None, arr, self.n, self.rank, byref(self.h)))

    def order(self, octant):
        out = (c_int32 * self.n)()
        check(lib.vr_sortlast_order(self.h, int(octant), out))
        return list(out)

    def close(self):
        if self.h:
            lib.vr_sortlast_destroy(self.h)
            self.h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------------
# device ops: the C ABI on torch CUDA tensors (torch is plumbing: memory + streams + collectives)
class CudaSlabOps:
    def __init__(self, ctx, sortlast, volume, width, height, count_samples=False):
        import torch
        self.torch, self.ctx, self.sl, self.vol = torch, ctx, sortlast, volume
        self.W, self.H = int(width), int(height)
        dev = torch.device("cuda", ctx.device)
        f32 = dict(dtype=torch.float32, device=dev)
        self.cseg = torch.empty((self.H, self.W), **f32)
        self.tseg = torch.empty((self.H, self.W), **f32)
        self.contrib = torch.empty((self.H, self.W, 2), **f32)
        self.rgba = torch.empty((self.H, self.W, 4), **f32)
        self.hit = torch.empty((self.H, self.W), dtype=torch.int32, device=dev)
        self.samples = torch.zeros(1, dtype=torch.int64, device=dev) if count_samples else None

    def _cnt(self):
        return c_void_p(self.samples.data_ptr()) if self.samples is not None else None

    def partial_composite(self, p):
        check(lib.vr_partial_composite(self.ctx.h, self.sl.h, self.vol.h, byref(p), c_void_p(self.cseg.data_ptr()),
                                       c_void_p(self.tseg.data_ptr()), self._cnt()))
        return self.cseg, self.tseg

    def resolve_composite(self, p, tseg_all, cseg):
        assert tseg_all.is_contiguous() and tseg_all.shape[0] == self.sl.n
        check(lib.vr_resolve_composite(self.ctx.h, self.sl.h, self.vol.h, byref(p), c_void_p(tseg_all.data_ptr()),
                                       c_void_p(cseg.data_ptr()), c_void_p(self.contrib.data_ptr()), self._cnt()))
        return self.contrib

    def resolve_composite_tin(self, p, tin_ptr, tseg, cseg):
        check(lib.vr_resolve_composite_tin(self.ctx.h, self.sl.h, self.vol.h, byref(p), c_void_p(int(tin_ptr)), c_void_p(tseg.data_ptr()),
                                           c_void_p(cseg.data_ptr()), c_void_p(self.contrib.data_ptr()), self._cnt()))
        return self.contrib

    def finish_composite(self, p, total):
        check(lib.vr_finish_composite(self.ctx.h, byref(p), c_void_p(total.data_ptr()), c_void_p(self.rgba.data_ptr())))
        return self.rgba

    def partial_iso(self, p):
        check(lib.vr_partial_iso(self.ctx.h, self.sl.h, self.vol.h, byref(p), c_void_p(self.rgba.data_ptr()),
                                 c_void_p(self.hit.data_ptr()), self._cnt()))
        return self.rgba, self.hit

    def resolve_iso(self, p, hit, hit_min, rgba):
        check(lib.vr_resolve_iso(self.ctx.h, self.sl.h, byref(p), c_void_p(hit.data_ptr()), c_void_p(hit_min.data_ptr()),
                                 c_void_p(rgba.data_ptr())))
        return rgba

    # halo boxes <-> dense buffers
    def pack_box(self, origin, extent, pad16=False):
        if pad16:  # flat buffer rounded up to 16 bytes (peer puts move 16-byte words)
            n = int(extent[0]) * int(extent[1]) * int(extent[2])
            buf = self.torch.empty(((n + 15) // 16 * 16,), dtype=self.torch.uint8, device=self.cseg.device)
        else:
            buf = self.torch.empty((extent[2], extent[1], extent[0]), dtype=self.torch.uint8, device=self.cseg.device)
        check(lib.vr_volume_pack_box(self.ctx.h, self.vol.h, (c_int64 * 3)(*origin), (c_int64 * 3)(*extent), c_void_p(buf.data_ptr())))
        return buf

    def unpack_box_ptr(self, origin, extent, dev_ptr):
        check(lib.vr_volume_unpack_box(self.ctx.h, self.vol.h, (c_int64 * 3)(*origin), (c_int64 * 3)(*extent), c_void_p(int(dev_ptr))))

    def unpack_box(self, origin, extent, buf):
        assert buf.is_contiguous()
        check(lib.vr_volume_unpack_box(self.ctx.h, self.vol.h, (c_int64 * 3)(*origin), (c_int64 * 3)(*extent), c_void_p(buf.data_ptr())))

    def empty_box(self, extent):
        return self.torch.empty((extent[2], extent[1], extent[0]), dtype=self.torch.uint8, device=self.cseg.device)


# ---------------------------------------------------------------------------------------------------
# exchange
class TorchDistExchange:
    """The frame's collectives over torch.distributed (NCCL over NVLink on GPUs, gloo in CPU tests)."""

    def __init__(self, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.n = dist.get_world_size(group)
        self._gather = None

    def all_gather(self, t):
        if self._gather is None or self._gather.shape[1:] != t.shape or self._gather.device != t.device:
            self._gather = self.torch.empty((self.n,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        self.dist.all_gather_into_tensor(self._gather.view(-1), t.contiguous().view(-1), group=self.group)
        return self._gather

    def all_reduce_sum(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_min(self, t):
        out = t.clone()
        self.dist.all_reduce(out, op=self.dist.ReduceOp.MIN, group=self.group)
        return out


class _DevMem:
    """Device memory owned by the library, exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerExchange:
    """The frame's exchange steps over peer memory (vr_peer_*, include/vr_b200.h): every rank owns a window that all
    ranks of the node map; a step is ONE store kernel over NVLink plus a flag, the consumer waits on the device.
    Same interface as TorchDistExchange (tensors in, tensors out), plus exchange_halos().

    Window layout for P pixels and n ranks (bytes):  gather n*4P | stripes-in n*(16P/n) | total 16P | halo staging.
    handles: callable(bytes64) -> list of the n ranks' 64-byte handles (e.g. an all-gather over torch.distributed);
    None = raw mode, call open_raw() with the windows of same-process ranks."""

    GATHER, RED_IN, RED_OUT, HALO = 0, 1, 2, 3

    def __init__(self, ctx, rank, n, W, H, halo_bytes=0, handles=None, halo_buffers=1):
        import torch
        self.torch, self.ctx, self.rank, self.n, self.P = torch, ctx, int(rank), int(n), int(W) * int(H)
        P = self.P
        assert (4 * P) % (16 * n) == 0, "image size must split into 16-byte aligned stripes"
        self.off_gather = 0
        self.off_red_in = self.off_gather + 4 * P * n
        self.off_total = self.off_red_in + 16 * P          # n stripes of 16P/n bytes (largest payload: rgba float4)
        self.off_halo = self.off_total + 16 * P
        self.halo_bytes = (int(halo_bytes) + 15) // 16 * 16
        self.halo_buffers = int(halo_buffers)  # staging areas used in turn (InterleavedSteps alternates two)
        self.bytes = self.off_halo + self.halo_bytes * self.halo_buffers
        self.h = c_void_p()
        check(lib.vr_peer_create(ctx.h, self.rank, self.n, self.bytes, byref(self.h)))
        self.device = torch.device("cuda", ctx.device if hasattr(ctx, "device") else torch.cuda.current_device())
        self._scratch = {}
        if handles is not None:
            mine = (ctypes.c_uint8 * 64)()
            check(lib.vr_peer_handle(self.h, mine))
            allh = handles(bytes(mine))
            blob = (ctypes.c_uint8 * (64 * self.n)).from_buffer_copy(b"".join(allh))
            check(lib.vr_peer_open(self.h, blob))
        self.base = None

    def raw(self):
        w = c_void_p()
        check(lib.vr_peer_raw(self.h, byref(w)))
        return w.value

    def open_raw(self, windows):
        check(lib.vr_peer_open_raw(self.h, (c_void_p * self.n)(*windows)))

    def _base(self):
        if self.base is None:
            w = c_void_p()
            check(lib.vr_peer_window(self.h, self.rank, byref(w)))
            self.base = w.value
        return self.base

    def _view(self, off, shape, dtype):
        ts = {self.torch.float32: "<f4", self.torch.int32: "<i4", self.torch.uint8: "|u1"}[dtype]
        return self.torch.as_tensor(_DevMem(self._base() + off, shape, ts), device=self.device)

    def _put(self, slot, copies):
        k = len(copies)
        check(lib.vr_peer_put(self.h, slot, k, (c_int32 * k)(*[c[0] for c in copies]), (c_int64 * k)(*[c[1] for c in copies]),
                              (c_void_p * k)(*[c[2] for c in copies]), (c_int64 * k)(*[c[3] for c in copies])))

    def _wait(self, slot, ranks):
        check(lib.vr_peer_wait(self.h, slot, len(ranks), (c_int32 * len(ranks))(*ranks)))

    # Every collective is a short sequence of stages; a stage never depends on a peer's work of the same stage, so
    # same-process ranks can be driven stage by stage (render_frame_peer_lockstep) while one-process-per-GPU ranks
    # simply run them back to back.
    def stages(self, kind, t):
        """Generator: runs one stage per next(); the last value is the collective's result."""
        assert t.is_contiguous() and t.element_size() == 4
        if kind == "all_gather":
            assert t.numel() == self.P
            nb = 4 * self.P
            self._put(self.GATHER, [(r, self.off_gather + self.rank * nb, t.data_ptr(), nb) for r in range(self.n)])
            yield None
            self._wait(self.GATHER, list(range(self.n)))
            yield self._view(self.off_gather, (self.n,) + tuple(t.shape), t.dtype)
            return
        # reduce-scatter + all-gather over the windows: rank j combines stripe j of every rank (fixed order), then
        # stores the finished stripe into every window
        op = {"all_reduce_sum": 0, "all_reduce_min": 1}[kind]
        nbytes = t.numel() * 4
        assert nbytes <= 16 * self.P and nbytes % (16 * self.n) == 0
        sb = nbytes // self.n
        base = t.data_ptr()
        self._put(self.RED_IN, [(j, self.off_red_in + self.rank * sb, base + j * sb, sb) for j in range(self.n)])
        yield None
        self._wait(self.RED_IN, list(range(self.n)))
        key = (sb, t.dtype)
        if key not in self._scratch:
            self._scratch[key] = self.torch.empty(sb // 4, dtype=t.dtype, device=t.device)
        red = self._scratch[key]
        ins = (c_void_p * self.n)(*[self._base() + self.off_red_in + k * sb for k in range(self.n)])
        check(lib.vr_peer_reduce(self.h, op, self.n, ins, c_void_p(red.data_ptr()), sb // 4))
        self._put(self.RED_OUT, [(r, self.off_total + self.rank * sb, red.data_ptr(), sb) for r in range(self.n)])
        yield None
        self._wait(self.RED_OUT, list(range(self.n)))
        yield self._view(self.off_total, tuple(t.shape), t.dtype)

    def _run(self, kind, t):
        res = None
        for res in self.stages(kind, t):
            pass
        return res

    def all_gather(self, t):
        return self._run("all_gather", t)

    def all_reduce_sum(self, t):
        return self._run("all_reduce_sum", t)

    def all_reduce_min(self, t):
        return self._run("all_reduce_min", t)

    def frame_stages(self, ops, params):
        """One COMPOSITE frame with the exchange fused into the march kernels (vr_peer_*_composite): three compute
        kernels that store into the peers' windows, three device-side waits.  Generator like stages(); the last value
        is the finished frame [H, W, 4] (a view of this rank's window)."""
        assert params.mode == COMPOSITE
        everyone = list(range(self.n))
        check(lib.vr_peer_partial_composite(self.ctx.h, self.h, self.GATHER, ops.sl.h, ops.vol.h, byref(params),
                                            c_void_p(ops.cseg.data_ptr()), self.off_gather, ops._cnt()))
        yield None
        self._wait(self.GATHER, everyone)
        check(lib.vr_peer_resolve_composite(self.ctx.h, self.h, self.RED_IN, ops.sl.h, ops.vol.h, byref(params), self.off_gather,
                                            c_void_p(ops.cseg.data_ptr()), self.off_red_in, ops._cnt()))
        yield None
        self._wait(self.RED_IN, everyone)
        check(lib.vr_peer_finish_composite(self.ctx.h, self.h, self.RED_OUT, byref(params), self.off_red_in, self.off_total))
        yield None
        self._wait(self.RED_OUT, everyone)
        yield self._view(self.off_total, (params.height, params.width, 4), self.torch.float32)

    TIN = 4
    T_IN_MIN_RANKS = 5  # from this many ranks up the T_in stripe exchange replaces the all-gather of T_seg (measured: slower on 2,
                        # equal on 4 -- 0.511 vs 0.509 ms per step -- faster on 8 GPUs)

    def frame_stages_hybrid(self, ops, params, frame_to=None, t_in=None):
        """COMPOSITE frame, the default: plain march kernels, the hand-overs as coalesced copy kernels into the peers' windows,
        the last pass (sum of the ranks' terms + colour map) as ONE kernel that stores the finished stripes.  Five stages.
        t_in (default: n > 4): instead of an all-gather of T_seg (n - 1 image-sized transfers per rank), stripe j of T_seg goes
        to rank j, which hands every rank the transmittance in front of its segment (vr_peer_tin_composite): 2 (n-1)/n
        transfers and one more round trip -- measured slower on 2 GPUs (0.30 vs 0.28 ms per frame), faster on 8.  T_in lives
        in the unused half of the stripes-in area (composite terms need 8P of its 16P bytes).
        frame_to = r: the finished frame goes to rank r only (the last value is None on the others)."""
        assert params.mode == COMPOSITE
        everyone = list(range(self.n))
        if t_in is None:
            t_in = self.n >= self.T_IN_MIN_RANKS
        cseg, tseg = ops.partial_composite(params)
        if t_in:
            s4 = 4 * self.P // self.n
            self._put(self.GATHER, [(j, self.off_gather + self.rank * s4, tseg.data_ptr() + j * s4, s4) for j in everyone])
            yield None
            self._wait(self.GATHER, everyone)
            off_tin = self.off_red_in + 8 * self.P
            check(lib.vr_peer_tin_composite(self.ctx.h, self.h, self.TIN, ops.sl.h, byref(params), self.off_gather, off_tin))
            yield None
            self._wait(self.TIN, everyone)
            contrib = ops.resolve_composite_tin(params, self._base() + off_tin, tseg, cseg)
        else:
            nb = 4 * self.P
            self._put(self.GATHER, [(r, self.off_gather + self.rank * nb, tseg.data_ptr(), nb) for r in everyone])
            yield None
            yield None
            self._wait(self.GATHER, everyone)
            contrib = ops.resolve_composite(params, self._view(self.off_gather, (self.n,) + tuple(tseg.shape), tseg.dtype), cseg)
        sb = 8 * self.P // self.n
        self._put(self.RED_IN, [(j, self.off_red_in + self.rank * sb, contrib.data_ptr() + j * sb, sb) for j in everyone])
        yield None
        self._wait(self.RED_IN, everyone)
        check(lib.vr_peer_finish_composite_to(self.ctx.h, self.h, self.RED_OUT, byref(params), self.off_red_in, self.off_total,
                                              -1 if frame_to is None else int(frame_to)))
        yield None
        if frame_to is not None and frame_to != self.rank:
            yield None
            return
        self._wait(self.RED_OUT, everyone)
        yield self._view(self.off_total, (params.height, params.width, 4), self.torch.float32)

    def render_frame(self, ops, params, fused=False, frame_to=None):
        """The frame of one rank (one process per GPU).  fused=True (COMPOSITE only) lets the march kernels store into
        the peers' windows themselves; measured slower than the staged copies on 2 x B200 (0.334 vs 0.287 ms per
        1024^2 frame: 32-byte stores in the 8x4 pixel-tile order and a system-scope fence per CTA inside the compute
        kernels cost more than the separate coalesced copy kernel), so the staged exchange is the default."""
        if params.mode != COMPOSITE:
            return render_frame(ops, self, params)
        res = None
        for res in (self.frame_stages(ops, params) if fused else self.frame_stages_hybrid(ops, params, frame_to=frame_to)):
            pass
        return res

    def exchange_halos(self, ops, descs, stored, defer_receive=False, buf=0):
        """Halo boxes: the owner packs a box and stores it into the staging area of the rank that needs it; that rank
        waits on the device and unpacks into its slab (which also refreshes the occupancy of the box).
        buf: which of the halo_buffers staging areas this exchange uses (same on all ranks)."""
        assert 0 <= buf < self.halo_buffers
        base = self.off_halo + buf * self.halo_bytes
        rank, my_o = self.rank, stored[self.rank][0]
        offs = {}            # running staging offset per destination, identical on all ranks
        copies, keep, recvs = [], [], []
        for src, dst, lo, ext in halo_transfers(descs, stored):
            nb = (int(ext[0]) * int(ext[1]) * int(ext[2]) + 15) // 16 * 16
            o = offs.get(dst, 0)
            offs[dst] = o + nb
            assert o + nb <= self.halo_bytes, "halo staging area too small"
            if src == rank:
                pk = ops.pack_box([lo[a] - stored[src][0][a] for a in range(3)], ext, pad16=True)
                keep.append(pk)
                copies.append((dst, base + o, pk.data_ptr(), nb))
            elif dst == rank:
                recvs.append((src, [lo[a] - my_o[a] for a in range(3)], ext, base + o))
        if copies:
            self._put(self.HALO, copies)
        self._keep = getattr(self, "_keep_prev", None), keep  # the put kernel reads them asynchronously: held for two calls
        self._keep_prev = keep
        if defer_receive:
            return recvs
        self.receive_halos(ops, recvs)
        return None

    def receive_halos(self, ops, recvs):
        if recvs:
            self._wait(self.HALO, sorted({r[0] for r in recvs}))
            for _, o3, ext, o in recvs:
                ops.unpack_box_ptr(o3, ext, self._base() + o)

    def timed_out(self):
        e = c_int32(0)
        check(lib.vr_peer_error(self.h, byref(e)))
        self.last_error = e.value
        return bool(e.value)

    def close(self):
        if self.h:
            lib.vr_peer_destroy(self.h)
            self.h = c_void_p()


class SortLastFrame:
    """vr_sortlast_frame: the whole frame / the halo exchange of a timestep as ONE C call each (csrc/vr_capi_frame.cu) -- what
    a C++ host uses instead of PeerExchange.  The peer window must be at least window_bytes() large."""

    @staticmethod
    def window_bytes(descs, W, H, halo, halo_buffers=1):
        n = len(descs)
        out = c_int64()
        check(lib.vr_sortlast_frame_window_bytes((SlabDesc * n)(*descs), n, int(W), int(H), (c_int64 * 3)(*[int(v) for v in halo]),
                                                 int(halo_buffers), byref(out)))
        return out.value

    def __init__(self, ctx, peer_handle, sl, W, H, halo, halo_buffers=1):
        import torch
        self.torch, self.ctx, self.W, self.H = torch, ctx, int(W), int(H)
        self.h = c_void_p()
        check(lib.vr_sortlast_frame_create(ctx.h, peer_handle, sl.h, self.W, self.H, (c_int64 * 3)(*[int(v) for v in halo]),
                                           int(halo_buffers), byref(self.h)))
        self.device = torch.device("cuda", ctx.device if hasattr(ctx, "device") else torch.cuda.current_device())

    def exchange_halos(self, vol, buffer=0, part=0):
        check(lib.vr_sortlast_exchange_halos(self.h, vol.h, int(buffer), int(part)))

    def render(self, vol, params, samples_ptr=None):
        """Issues the frame; returns it as a tensor view [H, W, 4] of this rank's window."""
        out = c_void_p()
        check(lib.vr_sortlast_render(self.h, vol.h, byref(params), byref(out), c_void_p(samples_ptr) if samples_ptr else None))
        return self.torch.as_tensor(_DevMem(out.value, (self.H, self.W, 4), "<f4"), device=self.device)

    def close(self):
        if self.h:
            lib.vr_sortlast_frame_destroy(self.h)
            self.h = c_void_p()


class InterleavedSteps:
    """Steady state of a time-varying sharded volume on ONE stream (N > 1, peer-memory exchange): step t draws the COMPOSITE
    frame of the slab decoded by step t-1 and decodes the next timestep into the other slab, with the stages of the two
    interleaved so that every device-side wait sits behind a long kernel of the other stage:

        partial(t) + put T_seg stripes  |  wait frame(t-1)  |  DECODE(t+1)  |  wait stripes, T_in to every rank  |  halo pack + put (t+1)
        |  wait T_in, resolve(t), put terms  |  wait halo, unpack (t+1)  |  wait terms, sum + colour map + broadcast (t)
                                                                                           [frame(t) is waited for in step t+1]

    The first exchange's latency and the skew between ranks' pass 1 hide behind the decode, the T_in round trip behind the halo
    pack, the reduce-scatter's behind the halo unpack, the broadcast's behind the next step's pass 1.  No second stream, no concurrency between kernels is needed (a
    persistent decode launch leaves no room for it).  Two halo staging areas alternate: a neighbour can be a whole step ahead.
    Write-after-read on the windows: a rank's next store into an area follows a wait that the area's reader preceded in its
    own stream order (T_seg stripes: put(t+1) follows wait terms(t), whose signal follows the owner's T_in pass (t); T_in: the owner's
    pass (t+1) follows its wait stripes(t+1), whose signal follows the reader's resolve(t); terms: put(t+1) follows wait T_in(t+1),
    whose signal follows the owner's wait stripes(t+1) and so its sum(t); frame: the sum of t+1 follows wait terms(t+1), whose
    signal follows the reader's wait frame(t); halo area b: put(t+2) follows wait terms(t), whose signal follows unpack(t))."""

    def __init__(self, peer, plans, ops, descs, stored, params, frame_to=None):
        assert peer.halo_buffers >= 2 and len(plans) == 2 and len(ops) == 2 and params.mode == COMPOSITE
        self.peer, self.plans, self.ops, self.descs, self.stored, self.p = peer, plans, ops, descs, stored, params
        self.frame_to = frame_to  # None: every rank receives the finished frame; r: rank r only
        self.pending = None   # generator of the frame whose broadcast has not been waited for
        self.k = 0            # slab the next step draws
        self.t = 0

    def prologue(self):
        """Decode slab 0 and complete its halo: the state every step expects on entry."""
        self.plans[0].run()
        self.peer.exchange_halos(self.ops[0], self.descs, self.stored, buf=0)
        self.k, self.t = 0, 1

    def step(self, consume=None):
        """One step: frame of slab k (issued, not yet waited for) + decode and halo of slab 1-k.  consume(frame): called with
        the previous step's finished frame (a view of this rank's window) as soon as it has been waited for; whatever reads
        it must be enqueued on the context's stream inside the callback -- later the peers may overwrite it."""
        k, peer = self.k, self.peer
        gen = peer.frame_stages_hybrid(self.ops[k], self.p, frame_to=self.frame_to)
        next(gen)                                         # pass 1 + T_seg stripes to their owners
        prev = next(self.pending) if self.pending is not None else None   # wait for the previous frame's stripes
        if consume is not None and prev is not None:
            consume(prev)
        self.plans[1 - k].run()                           # decode of the next timestep
        next(gen)                                         # wait T_seg stripes, T_in of my stripe to every rank
        recvs = peer.exchange_halos(self.ops[1 - k], self.descs, self.stored, defer_receive=True, buf=self.t & 1)
        next(gen)                                         # wait T_in, resolve, store terms into the stripe owners
        peer.receive_halos(self.ops[1 - k], recvs)        # wait halo planes, unpack (+ occupancy of the box)
        next(gen)                                         # wait terms, sum + colour map + stripes into every window
        self.pending = gen
        self.k, self.t = 1 - k, self.t + 1

    def drain(self):
        """Wait for the last issued frame; returns it."""
        out = next(self.pending) if self.pending is not None else None
        self.pending = None
        return out


def frame_steps(ops, params):
    """One sort-last frame of one rank as a coroutine: yields (collective, tensor) requests, receives the
    collective's result, returns the finished RGBA frame [H, W, 4] (row 0 = bottom row)."""
    if params.mode == COMPOSITE:
        cseg, tseg = ops.partial_composite(params)
        tseg_all = yield ("all_gather", tseg)
        contrib = ops.resolve_composite(params, tseg_all, cseg)
        total = yield ("all_reduce_sum", contrib)
        return ops.finish_composite(params, total)
    rgba, hit = ops.partial_iso(params)
    hit_min = yield ("all_reduce_min", hit)
    rgba = ops.resolve_iso(params, hit, hit_min, rgba)
    total = yield ("all_reduce_sum", rgba)
    return total


def render_frame(ops, exchange, params):
    """Drive frame_steps with real collectives (one process per GPU)."""
    gen = frame_steps(ops, params)
    try:
        req = next(gen)
        while True:
            kind, t = req
            req = gen.send(getattr(exchange, kind)(t))
    except StopIteration as fin:
        return fin.value


def render_frame_peer_lockstep(ops_list, ex_list, params):
    """Virtual ranks of ONE process over PeerExchange windows: every collective is driven stage by stage over all
    ranks, so no rank's device-side wait is queued before the put it waits for."""
    gens = [frame_steps(o, params) for o in ops_list]
    reqs = [next(g) for g in gens]
    done = [None] * len(gens)
    while any(d is None for d in done):
        kind = reqs[0][0]
        assert all(r[0] == kind for r in reqs)
        st = [ex.stages(kind, r[1]) for ex, r in zip(ex_list, reqs)]
        res = [None] * len(st)
        alive = True
        while alive:
            alive = False
            for i, g in enumerate(st):
                try:
                    res[i] = next(g)
                    alive = True
                except StopIteration:
                    pass
        for i, g in enumerate(gens):
            try:
                reqs[i] = g.send(res[i])
            except StopIteration as fin:
                done[i] = fin.value
    return done


def render_frame_fused_lockstep(ops_list, ex_list, params, hybrid=False):
    """Virtual ranks of one process: PeerExchange.frame_stages (or frame_stages_hybrid) of all ranks, stage by stage."""
    st = [(ex.frame_stages_hybrid if hybrid else ex.frame_stages)(o, params) for ex, o in zip(ex_list, ops_list)]
    res = [None] * len(st)
    alive = True
    while alive:
        alive = False
        for i, g in enumerate(st):
            try:
                res[i] = next(g)
                alive = True
            except StopIteration:
                pass
    return res


def render_frame_lockstep(ops_list, params):
    """All ranks in ONE process (virtual ranks on one device, or CPU stand-ins): the collectives are done
    locally on the list of tensors.  Returns every rank's finished frame."""
    import torch
    gens = [frame_steps(o, params) for o in ops_list]
    reqs = [next(g) for g in gens]
    done = [None] * len(gens)
    while any(d is None for d in done):
        kind = reqs[0][0]
        assert all(r[0] == kind for r in reqs)
        ts = [r[1] for r in reqs]
        if ts[0].is_cuda:  # the ranks' kernels may run on streams of their own; the local collective runs on torch's
            torch.cuda.synchronize()
        if kind == "all_gather":
            res = [torch.stack(ts).contiguous()] * len(ts)
        elif kind == "all_reduce_sum":
            tot = torch.stack(ts).sum(0)
            res = [tot.clone() for _ in ts]
        elif kind == "all_reduce_min":
            tot = torch.stack(ts).amin(0)
            res = [tot.clone() for _ in ts]
        else:
            raise ValueError(kind)
        if ts[0].is_cuda:
            torch.cuda.synchronize()
        for i, g in enumerate(gens):
            try:
                reqs[i] = g.send(res[i])
            except StopIteration as fin:
                done[i] = fin.value
    return done


# ---------------------------------------------------------------------------------------------------
# halo exchange (once per decoded timestep, not per frame)
def exchange_halos(ops, descs, stored, rank, dist=None, group=None):
    """Fill rank's halo from the ranks that own it and serve the others' halos (grouped point-to-point
    sends: ncclSend/ncclRecv under torch.distributed.batch_isend_irecv)."""
    import torch.distributed as tdist
    dist = dist or tdist
    my_o = stored[rank][0]
    p2p, recvs = [], []
    for src, dst, lo, ext in halo_transfers(descs, stored):
        if src == rank:
            o_src = [lo[a] - stored[src][0][a] for a in range(3)]
            p2p.append(dist.P2POp(dist.isend, ops.pack_box(o_src, ext), dst, group=group))
        elif dst == rank:
            buf = ops.empty_box(ext)
            recvs.append(([lo[a] - my_o[a] for a in range(3)], ext, buf))
            p2p.append(dist.P2POp(dist.irecv, buf, src, group=group))
    if p2p:
        for w in dist.batch_isend_irecv(p2p):
            w.wait()
    for o, ext, buf in recvs:
        ops.unpack_box(o, ext, buf)


def exchange_halos_lockstep(ops_list, descs, stored):
    """Same for virtual ranks living in one process."""
    for src, dst, lo, ext in halo_transfers(descs, stored):
        buf = ops_list[src].pack_box([lo[a] - stored[src][0][a] for a in range(3)], ext)
        ops_list[dst].unpack_box([lo[a] - stored[dst][0][a] for a in range(3)], ext, buf)
