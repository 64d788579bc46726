"""CPU: the Python mirror of the reference's VolumeReader (VolumeReader.h:36-291) on its host path (toGPU=False):
placement of LoadBricksToTexture (:184-205) with the brick map of fillVolumeBrickMap (main.cpp:599-619), a missing file
returns False, a wrong file size raises (:258-260).  The toGPU path is vr_volume_load_bricks (tests/test_gpu_io.py)."""
import numpy as np
import pytest

import volumerenderer_b200 as vr
from volumerenderer_b200 import synth


def _scene(d, grid=(2, 3, 2), brick=(32, 16, 8), t=5):
    I, J, K = grid
    X, Y, Z = brick
    full = synth.noise_brick((I * X, J * Y, K * Z), seed=9)
    bmap = {}
    for b in range(I * J * K):
        i, j, k = b % I, (b // I) % J, b // (I * J)
        bmap[b] = (i, j, k)
        full[k * Z:(k + 1) * Z, j * Y:(j + 1) * Y, i * X:(i + 1) * X].tofile(str(d / f"brick{b}_{t}.raw"))
    return full, bmap


def test_load_bricks_host_path(tmp_path):
    full, bmap = _scene(tmp_path)
    r = vr.VolumeReader((32, 16, 8), (2, 3, 2), lambda b, t: tmp_path / f"brick{b}_{t}.raw", bmap)
    assert r.LoadBricksToTexture(12, 2, 3, 2, 5, False, toGPU=False)
    assert r.dataDims == (64, 48, 16) and np.array_equal(r.data, full)
    assert r.LoadBrickToTexture(7, 5, False, toGPU=False)
    i, j, k = bmap[7]
    assert np.array_equal(r.data, full[k * 8:(k + 1) * 8, j * 16:(j + 1) * 16, i * 32:(i + 1) * 32])


def test_missing_and_malformed_files(tmp_path):
    full, bmap = _scene(tmp_path)
    find = lambda b, t: tmp_path / f"brick{b}_{t}.raw"
    r = vr.VolumeReader((32, 16, 8), (2, 3, 2), find, bmap)
    assert r.LoadBricksToTexture(12, 2, 3, 2, 6, False, toGPU=False) is False   # no files for timestep 6
    assert r.LoadBrickToTexture(99, 5, False, toGPU=False) is False
    (tmp_path / "brick3_5.raw").write_bytes(b"\0" * 17)
    with pytest.raises(ValueError):
        r.LoadBricksToTexture(12, 2, 3, 2, 5, False, toGPU=False)
    del bmap[4]
    with pytest.raises(KeyError):
        r.LoadBricksToTexture(12, 2, 3, 2, 5, False, toGPU=False)
    with pytest.raises(vr.VrError):  # no CPU fallback behind transferToGPU
        vr.VolumeReader((32, 16, 8), (2, 3, 2), find, bmap).transferToGPU()
