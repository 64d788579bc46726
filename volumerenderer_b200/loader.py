"""Host-side mirror of the reference's brick loader, `template <typename T> class VolumeReader`
(/root/reference/volume_renderer/VolumeReader.h:36-291), with the reference's method names and argument meaning:

    VolumeReader(brickDims, gridDims, findSourceFile, brickMap)          VolumeReader.h:68
    LoadBrickToTexture(brick, timestep, dealloc, toGPU=True)             :91
    LoadBricksToTexture(numBricks, I, J, K, timestep, dealloc, toGPU)    :151
    transferToGPU(dealloc=True)                                           :114
    data, dataDims, volume (replaces textureId)                           :42-51

A missing file makes the Load* methods return False (as the reference does), a file whose size does not match the
brick dimensions raises (the reference throws, :258-260).  With toGPU the bricks go straight into a device slab through
vr_volume_load_bricks (pipelined reads and copies, 64-bit placement); without it the array is assembled on the host with
the placement of VolumeReader.h:184-205.  There is no CPU renderer behind this class: `volume` only exists on a GPU."""
import os

import numpy as np

from . import _lib
from .kdtree import Volume


class VolumeReader:
    def __init__(self, brickDims, gridDims, findSourceFile, brickMap, ctx=None):
        self.brickDims = tuple(int(v) for v in brickDims)   # X, Y, Z of one brick
        self.gridDims = tuple(int(v) for v in gridDims)
        self.findSourceFile = findSourceFile                 # (brick, timestep) -> path
        self.brickMap = brickMap                             # brick -> (i, j, k), fillVolumeBrickMap main.cpp:599-619
        self.ctx = ctx
        self.data = None                                     # uint8 (Z, Y, X), x fastest
        self.dataDims = (0, 0, 0)
        self.volume = None

    # LoadVolumeFromBinaryFile, VolumeReader.h:244-289
    def _read_brick(self, path):
        X, Y, Z = self.brickDims
        if not os.path.exists(path):
            return None
        if os.path.getsize(path) != X * Y * Z:
            raise ValueError(f"size of {path} does not match the brick dimensions")
        return np.fromfile(path, np.uint8).reshape(Z, Y, X)

    def LoadBrickToTexture(self, brick, timestep, dealloc, toGPU=True):
        b = self._read_brick(str(self.findSourceFile(brick, timestep)))
        if b is None:
            return False
        self.data, self.dataDims = b, self.brickDims
        if toGPU:
            self.transferToGPU(dealloc)
        return True

    def LoadBricksToTexture(self, numBricks, I, J, K, timestep, dealloc, toGPU=True):
        X, Y, Z = self.brickDims
        self.dataDims = (I * X, J * Y, K * Z)
        for b in range(numBricks):
            if b not in self.brickMap:
                raise KeyError(f"brick map has no entry for brick {b}")
        if toGPU:
            if self.ctx is None:
                raise _lib.VrError(_lib.VR_ERR_STATE, "LoadBricksToTexture(toGPU=True): no Context given")
            if self.volume is not None:
                self.volume.close()
            self.volume = Volume(self.ctx, *self.dataDims)
            try:
                res = self.volume.load_bricks(numBricks, I, J, K, timestep, self.brickDims, self.findSourceFile,
                                              brick_map=self.brickMap, want_host=not dealloc)
            except _lib.VrError as e:
                if e.code == _lib.VR_ERR_IO:
                    return False
                if e.code == _lib.VR_ERR_FORMAT:
                    raise ValueError(str(e)) from e
                raise
            self.data = None if dealloc else res[2]
            return True
        out = np.zeros((K * Z, J * Y, I * X), np.uint8)
        for b in range(numBricks):
            i, j, k = self.brickMap[b]
            brick = self._read_brick(str(self.findSourceFile(b, timestep)))
            if brick is None:
                return False
            out[k * Z:(k + 1) * Z, j * Y:(j + 1) * Y, i * X:(i + 1) * X] = brick  # VolumeReader.h:184-205
        self.data = out
        return True

    def transferToGPU(self, dealloc=True):
        if self.ctx is None:
            raise _lib.VrError(_lib.VR_ERR_STATE, "transferToGPU: no Context given")
        if self.data is None:
            raise _lib.VrError(_lib.VR_ERR_STATE, "transferToGPU: no data loaded")
        if self.volume is not None:
            self.volume.close()
        self.volume = Volume(self.ctx, *self.dataDims)
        self.volume.upload_brick(self.data, (0, 0, 0))
        if dealloc:
            self.data = None
