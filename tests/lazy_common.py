"""Shared by the lazy-decode tests (CPU and GPU): a small multi-brick volume with a void side, a solid side and a
noisy interface between them (the shape of the synthetic Richtmyer-Meshkov field, SURVEY 8d, shrunk)."""
import numpy as np

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import synth

BRICK = (64, 64, 32)  # X, Y, Z: tiles are 32x32x16 = one occupancy cell, 8 tiles per brick


def layered_volume(dims=(256, 256, 256), seed=7):
    """uint8 (Z, Y, X): 0 below a wavy interface, 230 above it, a noisy ramp of half-width 6 in between."""
    X, Y, Z = dims
    gx = np.arange(X, dtype=np.float64)[None, None, :]
    gy = np.arange(Y, dtype=np.float64)[None, :, None]
    gz = np.arange(Z, dtype=np.float64)[:, None, None]
    tw = 2.0 * np.pi / X
    H = Z * 0.5 + Z * 0.12 * (np.sin(tw * 2 * gx) * np.cos(tw * 1 * gy) + 0.5 * np.sin(tw * 5 * (gx + gy)))
    d = (gz - H) / 6.0
    ix = np.arange(X, dtype=np.uint32)[None, None, :]
    iy = np.arange(Y, dtype=np.uint32)[None, :, None]
    iz = np.arange(Z, dtype=np.uint32)[:, None, None]
    noise = ((synth.mix32(ix, iy, iz, 0, seed) >> np.uint32(24)) % np.uint32(33)).astype(np.float64) - 16.0
    mid = np.clip(np.round(115.0 * (1.0 + d)) + noise * (1.0 - np.abs(d)), 0, 255)
    return np.ascontiguousarray(np.where(d <= -1.0, 0.0, np.where(d >= 1.0, 230.0, mid)).astype(np.uint8))


def compress_bricks(vol, brick=BRICK, tol=1, epochs=2, compress=None):
    """Cuts vol (Z, Y, X) into bricks; returns (streams, origins [(x, y, z)])."""
    Z, Y, X = vol.shape
    bx, by, bz = brick
    streams, origins = [], []
    for k in range(Z // bz):
        for j in range(Y // by):
            for i in range(X // bx):
                b = vol[k * bz:(k + 1) * bz, j * by:(j + 1) * by, i * bx:(i + 1) * bx]
                s, _ = (compress or vr.Stream.compress)(b, tol, epochs)
                streams.append(s)
                origins.append((i * bx, j * by, k * bz))
    return streams, origins


def oracle_decode(streams, origins, dims, brick=BRICK):
    """levelCut of every stream (plain-C restatement) assembled like LoadBricksToTexture."""
    X, Y, Z = dims
    bx, by, bz = brick
    out = np.zeros((Z, Y, X), np.uint8)
    for s, (ox, oy, oz) in zip(streams, origins):
        hd = s.header
        rc, dec = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, bx, by, bz)
        assert rc == 0
        out[oz:oz + bz, oy:oy + by, ox:ox + bx] = dec
    return out
