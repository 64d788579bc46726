"""GPU parity (run with -m gpu on the B200): CUDA stage-1 decode through the C ABI vs the oracle's
levelCut restatement and the committed golden vectors of the reference, bit-exact."""
import glob
import os

import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


def _oracle(s):
    hd = s.header
    rc, out = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth,
                                 hd.X, hd.Y, hd.Z)
    assert rc == 0
    return out


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_levelcut_matches_reference_golden(ctx, path, tmp_path):
    g = np.load(path)
    p = tmp_path / "t.bin"
    p.write_bytes(g["container"].tobytes())
    t = vr.VolumeKdtree(ctx=ctx).open(p)
    out = t.levelCut(t.maxTreeDepth)
    assert np.array_equal(out, g["decoded"])


@pytest.mark.parametrize("dims", [(2, 2, 2), (8, 4, 2), (16, 16, 16), (64, 64, 32), (128, 64, 64), (16, 4, 64), (1024, 2, 2)])
@pytest.mark.parametrize("tol,ep", [(1, 2), (6, 5), (0, 5)])
def test_levelcut_matches_oracle_small(ctx, dims, tol, ep):
    X, Y, Z = dims
    for vol in (synth.shell_brick(dims, noise=6, seed=tol), synth.noise_brick(dims, seed=ep + 1)):
        t = vr.VolumeKdtree(vol, X, Y, Z, ctx=ctx)
        t.setErrorTolerance(tol); t.setMaxEpochs(ep)
        t.build()
        out = t.levelCut(t.maxTreeDepth)
        assert np.array_equal(out, _oracle(t.stream))


def test_levelcut_rejects_partial_cut(ctx):
    vol = synth.shell_brick((16, 16, 8))
    t = vr.VolumeKdtree(vol, 16, 16, 8, ctx=ctx).build()
    with pytest.raises(vr.VrError) as e:  # the reference segfaults here
        t.levelCut(t.maxTreeDepth - 1)
    assert e.value.code == _lib.VR_ERR_UNSUPPORTED


@pytest.mark.parametrize("kind", ["rm", "noise", "constant", "rm_empty"])
def test_full_size_brick_bit_exact(ctx, kind):
    """BASELINE config 1/2 brick shape, 256x256x128."""
    vol = {"rm": lambda: synth.rm_brick(3, 4, 7), "noise": synth.noise_brick, "constant": synth.constant_brick,
           "rm_empty": lambda: synth.rm_brick(0, 0, 0)}[kind]()
    t = vr.VolumeKdtree(vol, 256, 256, 128, ctx=ctx)
    t.setErrorTolerance(1); t.setMaxEpochs(2)
    t.build()
    out = t.levelCut(t.maxTreeDepth)
    assert np.array_equal(out, _oracle(t.stream))
    if kind != "noise":
        assert t.measureMaxError() <= 1


def test_batched_plan_places_bricks_like_LoadBricksToTexture(ctx):
    """Grid 2x2x2 of 32x32x16 bricks, brick b at (i,j,k) = (b%2,(b/2)%2,b/4) (main.cpp:599-619);
    slab larger than the grid, everything outside the bricks must stay untouched."""
    bx, by, bz = 32, 32, 16
    SX, SY, SZ = 96, 70, 40
    vol = vr.Volume(ctx, SX, SY, SZ)
    sentinel = np.full((SZ, SY, SX), 0x5A, np.uint8)
    vol.upload_brick(sentinel, (0, 0, 0))
    want = sentinel.copy()
    streams, origins = [], []
    for b in range(8):
        i, j, k = b % 2, (b // 2) % 2, b // 4
        brick = synth.rm_region(600 + i * bx, 900 + j * by, 940 + k * bz, bx, by, bz) if b != 5 else synth.constant_brick((bx, by, bz), 9)
        s, _ = vr.Stream.compress(brick, 1, 2)
        s.upload(ctx)
        streams.append(s)
        o = (16 + i * bx, 3 + j * by, 5 + k * bz)
        origins.append(o)
        want[o[2]:o[2] + bz, o[1]:o[1] + by, o[0]:o[0] + bx] = _oracle(s)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    plan.run()
    ctx.synchronize()
    assert np.array_equal(vol.download(), want)
    b = plan.bytes()
    assert b["voxels"] == 8 * bx * by * bz and b["stream"] > 0
    # idempotent: a second run gives the same slab
    plan.run(); ctx.synchronize()
    assert np.array_equal(vol.download(), want)


def test_unaligned_placement_uses_scalar_stores(ctx):
    vol = vr.Volume(ctx, 50, 40, 20)
    vol.upload_brick(np.zeros((20, 40, 50), np.uint8), (0, 0, 0))
    brick = synth.shell_brick((32, 32, 16), noise=4)
    s, _ = vr.Stream.compress(brick, 1, 2)
    s.upload(ctx)
    plan = vr.DecodePlan(ctx, [s], [(7, 5, 3)], vol)
    plan.run(); ctx.synchronize()
    got = vol.download()
    assert np.array_equal(got[3:19, 5:37, 7:39], _oracle(s))
    got[3:19, 5:37, 7:39] = 0
    assert not got.any()


@pytest.mark.parametrize("kernel", ["persistent", "persistent_tensor_stores", "persistent_lane_stores", "persistent_split", "classic"])
def test_slab_of_full_size_bricks_both_kernels(ctx, kernel):
    """A 2x1x3 grid of 256x256x128 bricks (empty, mixing layer x 3, noise, solid) through the persistent warp-specialised
    kernel (parse groups + TMA fill warp, kd_decode.cu) and through the classic one-CTA-per-tile kernel: both equal the
    levelCut restatement voxel for voxel, write nothing outside the bricks, and leave the same exact per-cell maxima."""
    vr.debug_set("decode_kernel", 0 if kernel.startswith("persistent") else 1)
    # how a parsed unit's voxels leave the SM: one tensor store per unit, or 16-byte stores of the lanes (default: chosen per plan)
    vr.debug_set("unit_store", {"persistent_tensor_stores": 1, "persistent_lane_stores": 2}.get(kernel, 0))
    vr.debug_set("decode_split", 1 if kernel == "persistent_split" else 0)  # constant groups as a kernel of their own, side stream
    try:
        bx, by, bz = 256, 256, 128
        vols = [synth.constant_brick(value=0), synth.rm_brick(3, 4, 7), synth.rm_brick(2, 4, 7), synth.noise_brick(),
                synth.rm_brick(3, 4, 8), synth.constant_brick(value=230)]
        SX, SY, SZ = 2 * bx + 32, by + 32, 3 * bz + 16
        vol = vr.Volume(ctx, SX, SY, SZ)
        sentinel = np.full((SZ, SY, SX), 0x5A, np.uint8)
        vol.upload_brick(sentinel, (0, 0, 0))
        want = sentinel.copy()
        streams, origins = [], []
        for b, v in enumerate(vols):
            s, _ = vr.Stream.compress(v, 1, 2)
            s.upload(ctx)
            streams.append(s)
            o = (32 + (b % 2) * bx, 32, 16 + (b // 2) * bz)
            origins.append(o)
            want[o[2]:o[2] + bz, o[1]:o[1] + by, o[0]:o[0] + bx] = _oracle(s)
        plan = vr.DecodePlan(ctx, streams, origins, vol)
        assert plan.kernel() == kernel.split("_")[0]
        for _ in range(2):  # idempotent
            plan.run()
            ctx.synchronize()
            got = vol.download()
            assert np.array_equal(got, want)
        occ = vol.occupancy(0)
        for b, o in enumerate(origins):  # cells of the decoded bricks: exact maxima of their 32x32x16 voxels
            blk = want[o[2]:o[2] + bz, o[1]:o[1] + by, o[0]:o[0] + bx].reshape(bz // 16, 16, by // 32, 32, bx // 32, 32).max(axis=(1, 3, 5))
            assert np.array_equal(occ[o[2] // 16:(o[2] + bz) // 16, o[1] // 32:(o[1] + by) // 32, o[0] // 32:(o[0] + bx) // 32], blk), b
        plan.close(); vol.close()
    finally:
        vr.debug_set("decode_kernel", 0)
        vr.debug_set("unit_store", 0)
        vr.debug_set("decode_split", 0)


def test_handles_may_be_released_in_any_order(ctx):
    """A host with a garbage collector frees the volume before the plan as often as after it: the plan is detached,
    refuses to run (VR_ERR_STATE) and can still be destroyed."""
    brick = synth.shell_brick((64, 64, 32), noise=4)
    s, _ = vr.Stream.compress_device(ctx, brick, 1, 2)
    s.upload(ctx)
    vol = vr.Volume(ctx, 64, 64, 32)
    plan, plan2 = vr.DecodePlan(ctx, [s], [(0, 0, 0)], vol), vr.DecodePlan(ctx, [s], [(0, 0, 0)], vol)
    plan.run(); ctx.synchronize()
    plan2.close()  # plan before volume
    vol.close()    # volume before plan
    with pytest.raises(_lib.VrError) as e:
        plan.run()
    assert e.value.code == _lib.VR_ERR_STATE
    plan.close()
    s.close()
