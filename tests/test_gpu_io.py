"""GPU (-m gpu): the file-facing entry points.  vr_volume_load_bricks = VolumeReader::LoadBricksToTexture
(VolumeReader.h:151-223) with pipelined reads and copies; vr_streams_load = one VolumeKdtree::open per brick
(main.cpp:276-284) on worker threads with the uploads overlapped; both checked against numpy placement / the oracle."""
import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import _lib, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = vr.Context(0)
    yield c
    c.close()


def _write_bricks(d, full, brick, grid, t, order=None):
    bx, by, bz = brick
    I, J, K = grid
    n = I * J * K
    bmap = {}
    for b in range(n):
        i, j, k = (b % I, (b // I) % J, b // (I * J)) if order is None else order[b]
        bmap[b] = (i, j, k)
        full[k * bz:(k + 1) * bz, j * by:(j + 1) * by, i * bx:(i + 1) * bx].tofile(str(d / f"b{b}_t{t}.raw"))
    return bmap


@pytest.mark.parametrize("brick,grid", [((32, 32, 16), (2, 2, 2)), ((64, 32, 8), (3, 2, 5)), ((256, 256, 128), (2, 1, 2))])
def test_load_bricks_places_files_like_LoadBricksToTexture(ctx, tmp_path, brick, grid):
    bx, by, bz = brick
    I, J, K = grid
    n = I * J * K
    full = synth.noise_brick((I * bx, J * by, K * bz), seed=11)
    rng = np.random.default_rng(3)
    perm = rng.permutation(n)
    order = [(int(q) % I, (int(q) // I) % J, int(q) // (I * J)) for q in perm]  # a scrambled brick map
    bmap = _write_bricks(tmp_path, full, brick, grid, 3, order)
    vol = vr.Volume(ctx, I * bx, J * by, K * bz)
    nb, nbytes, host = vol.load_bricks(n, I, J, K, 3, brick, lambda b, t: tmp_path / f"b{b}_t{t}.raw", brick_map=bmap, want_host=True)
    assert nb == n and nbytes == full.size
    assert np.array_equal(host, full)
    assert np.array_equal(vol.download(), full)
    # default map = fillVolumeBrickMap (main.cpp:599-619)
    _write_bricks(tmp_path, full, brick, grid, 4)
    vol2 = vr.Volume(ctx, I * bx, J * by, K * bz)
    assert vol2.load_bricks(n, I, J, K, 4, brick, lambda b, t: tmp_path / f"b{b}_t{t}.raw") == (n, full.size)
    assert np.array_equal(vol2.download(), full)
    vol.close(); vol2.close()


def test_load_bricks_errors(ctx, tmp_path):
    brick, grid = (32, 32, 16), (2, 1, 1)
    full = synth.noise_brick((64, 32, 16), seed=2)
    _write_bricks(tmp_path, full, brick, grid, 0)
    vol = vr.Volume(ctx, 64, 32, 16)
    with pytest.raises(_lib.VrError) as ei:  # no files for timestep 9: the reference returns false
        vol.load_bricks(2, 2, 1, 1, 9, brick, lambda b, t: tmp_path / f"b{b}_t{t}.raw")
    assert ei.value.code == _lib.VR_ERR_IO
    (tmp_path / "b1_t0.raw").write_bytes(b"\x00" * 100)  # wrong size: the reference throws (VolumeReader.h:258-260)
    with pytest.raises(_lib.VrError) as ei:
        vol.load_bricks(2, 2, 1, 1, 0, brick, lambda b, t: tmp_path / f"b{b}_t{t}.raw")
    assert ei.value.code == _lib.VR_ERR_FORMAT
    with pytest.raises(_lib.VrError) as ei:  # grid larger than the volume
        vol.load_bricks(4, 2, 2, 1, 0, brick, lambda b, t: tmp_path / f"b{b}_t{t}.raw")
    assert ei.value.code == _lib.VR_ERR_ARG
    vol.close()


def test_streams_load_uploads_and_decodes(ctx, tmp_path):
    bx, by, bz = 64, 64, 32
    paths, want, origins = [], [], []
    for b in range(6):
        i, k = b % 3, b // 3
        brick = synth.rm_region(640 + i * bx, 900, 930 + k * bz, bx, by, bz)
        s, _ = vr.Stream.compress(brick, 1, 2)
        p = tmp_path / f"tree{b}.bin"
        s.save(p)
        paths.append(p)
        hd = s.header
        rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, bx, by, bz)
        assert rc == 0
        want.append(dec)
        origins.append((i * bx, 0, k * bz))
    streams = vr.load_streams(paths, ctx, threads=3)
    vol = vr.Volume(ctx, 3 * bx, by, 2 * bz)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    plan.run()
    got = vol.download()
    for (ox, oy, oz), dec in zip(origins, want):
        assert np.array_equal(got[oz:oz + bz, oy:oy + by, ox:ox + bx], dec)
    plan.close(); vol.close()


def _six_bricks(tmp_path, indexed):
    bx, by, bz = 64, 64, 32
    paths, want, origins, streams = [], [], [], []
    for b in range(6):
        i, k = b % 3, b // 3
        brick = synth.rm_region(640 + i * bx, 900, 930 + k * bz, bx, by, bz) if b != 4 else synth.constant_brick((bx, by, bz), 61)
        s, _ = vr.Stream.compress(brick, 1, 2)
        p = tmp_path / (f"tree{b}.kdx" if indexed(b) else f"tree{b}.bin")
        (s.save_indexed if indexed(b) else s.save)(p)
        paths.append(p)
        hd = s.header
        rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, bx, by, bz)
        assert rc == 0
        want.append(dec); origins.append((i * bx, 0, k * bz)); streams.append(s)
    return (bx, by, bz), paths, want, origins, streams


def _decodes(ctx, dims, streams, origins, want):
    bx, by, bz = dims
    vol = vr.Volume(ctx, 3 * bx, by, 2 * bz)
    plan = vr.DecodePlan(ctx, streams, origins, vol)
    plan.run()
    got = vol.download()
    for (ox, oy, oz), dec in zip(origins, want):
        assert np.array_equal(got[oz:oz + bz, oy:oy + by, ox:ox + bx], dec)
    plan.close(); vol.close()


def test_streams_load_mixed_containers_repeatedly(ctx, tmp_path):
    """Plain and indexed containers in one load (one allocation + one H2D copy per call, the pinned staging buffer reused):
    three loads in a row, earlier sets still decodable after later loads (they own their device memory)."""
    dims, paths, want, origins, _ = _six_bricks(tmp_path, lambda b: b % 2 == 0)
    sets = [vr.load_streams(paths, ctx, threads=3) for _ in range(3)]
    for ss in sets:
        _decodes(ctx, dims, ss, origins, want)
    for s in sets[1]:
        s.close()
    _decodes(ctx, dims, sets[0], origins, want)
    # a loaded stream can still be saved / validated / moved to a fresh upload
    s = sets[2][1]
    assert s.validate() == s.header.numActiveNodes
    s.pin(); s.upload(ctx)
    _decodes(ctx, dims, sets[2], origins, want)


def test_stream_arena_one_copy_per_timestep(ctx, tmp_path):
    dims, paths, want, origins, streams = _six_bricks(tmp_path, lambda b: False)
    shared = streams + [streams[0]]  # a stream may appear twice in a plan (constant bricks share one)
    arena = vr.StreamArena(ctx, shared)
    assert arena.nbytes >= sum(sum(s.device_bytes()) for s in streams)
    arena.upload(ctx)
    _decodes(ctx, dims, streams, origins, want)
    arena.upload(ctx)  # again: every step of a streaming loop
    _decodes(ctx, dims, streams, origins, want)
    arena.close()      # the buffers live on with the streams
    _decodes(ctx, dims, streams, origins, want)
    streams[2].upload(ctx)  # a single stream can still be re-uploaded on its own (into its slice)
    _decodes(ctx, dims, streams, origins, want)
