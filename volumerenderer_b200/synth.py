"""Deterministic synthetic inputs of the reference's brick shape.

Richtmyer-Meshkov-shaped scalar field on the 2048x2048x1920 grid of the reference's data set
(brick 256x256x128, grid 8x8x15, brick number b = i + 8j + 64k: main.cpp:78-79,599-619), plus
the two extreme bricks (uniform noise, constant).  numpy only; uint32 wrap-around hashing.
"""
import numpy as np

BRICK = (256, 256, 128)      # X, Y, Z  (main.cpp:78)
GRID = (8, 8, 15)            # I, J, K  (main.cpp:79)
FULL = (2048, 2048, 1920)


def mix32(x, y, z, t, seed):
    x = np.asarray(x, np.uint32); y = np.asarray(y, np.uint32); z = np.asarray(z, np.uint32)
    with np.errstate(over="ignore"):
        h = (np.uint32(seed) ^ (x * np.uint32(0x9E3779B1)) ^ (y * np.uint32(0x85EBCA77))
             ^ (z * np.uint32(0xC2B2AE3D)) ^ np.uint32((int(t) * 0x27D4EB2F) & 0xFFFFFFFF))
        h = h ^ (h >> np.uint32(15)); h = h * np.uint32(0x2C1B3C6D)
        h = h ^ (h >> np.uint32(12)); h = h * np.uint32(0x297A2D39)
        h = h ^ (h >> np.uint32(15))
    return h


def rm_region(x0, y0, z0, nx, ny, nz, t=0):
    """uint8 array (nz, ny, nx) of the RM field at global origin (x0, y0, z0), timestep t."""
    gx = (x0 + np.arange(nx, dtype=np.float64))[None, None, :]
    gy = (y0 + np.arange(ny, dtype=np.float64))[None, :, None]
    gz = (z0 + np.arange(nz, dtype=np.float64))[:, None, None]
    tw = 2.0 * np.pi / 2048.0
    H = 960.0 + (60.0 + 40.0 * t) * (0.5 * np.sin(tw * 3 * gx) * np.cos(tw * 2 * gy)
                                     + 0.3 * np.sin(tw * 7 * gx + 1.3) * np.sin(tw * 5 * gy + 0.7)
                                     + 0.2 * np.sin(tw * 17 * (gx + gy)))
    w = 8.0 + 6.0 * t
    d = (gz - H) / w
    ix = (x0 + np.arange(nx, dtype=np.uint32))[None, None, :]
    iy = (y0 + np.arange(ny, dtype=np.uint32))[None, :, None]
    iz = (z0 + np.arange(nz, dtype=np.uint32))[:, None, None]
    noise = ((mix32(ix, iy, iz, t, 0xB200 + t) >> np.uint32(24)) % np.uint32(33)).astype(np.float64) - 16.0
    mid = np.clip(np.round(115.0 * (1.0 + d)) + noise * (1.0 - np.abs(d)), 0, 255)
    out = np.where(d <= -1.0, 0.0, np.where(d >= 1.0, 230.0, mid))
    return np.ascontiguousarray(out.astype(np.uint8))


def rm_brick(i, j, k, t=0, brick=BRICK):
    X, Y, Z = brick
    return rm_region(i * X, j * Y, k * Z, X, Y, Z, t)


def brick_ijk(b, grid=GRID):
    return b % grid[0], (b // grid[0]) % grid[1], b // (grid[0] * grid[1])


def noise_brick(brick=BRICK, seed=0xB200):
    X, Y, Z = brick
    ix = np.arange(X, dtype=np.uint32)[None, None, :]
    iy = np.arange(Y, dtype=np.uint32)[None, :, None]
    iz = np.arange(Z, dtype=np.uint32)[:, None, None]
    return np.ascontiguousarray((mix32(ix, iy, iz, 0, seed) >> np.uint32(24)).astype(np.uint8))


def constant_brick(brick=BRICK, value=77):
    X, Y, Z = brick
    return np.full((Z, Y, X), value, np.uint8)


def shell_brick(brick, noise=0, seed=1):
    """Small-case generator used by the parity tests: sine shells, zero outside a sphere."""
    X, Y, Z = brick
    z, y, x = np.meshgrid(np.arange(Z), np.arange(Y), np.arange(X), indexing="ij")
    r = np.sqrt((x - X / 2) ** 2 + (y - Y / 2) ** 2 + ((z - Z / 2) * X / max(Z, 1)) ** 2)
    v = 128 + 100 * np.sin(r * 0.35)
    if noise:
        v = v + np.random.default_rng(seed).integers(-noise, noise + 1, v.shape)
    v[r > X * 0.45] = 0
    return np.ascontiguousarray(np.clip(v, 0, 255).astype(np.uint8))


def rm_brick_constant_value(i, j, k, t=0, brick=BRICK):
    """0 / 230 when the whole brick lies below / above the mixing layer for every (x, y), else None
    (bound from |H - 960| <= 60 + 40 t and half-width w = 8 + 6 t)."""
    Z = brick[2]
    amp, w = 60.0 + 40.0 * t, 8.0 + 6.0 * t
    z0, z1 = k * Z, k * Z + Z - 1
    if z1 <= 960.0 - amp - w:
        return 0
    if z0 >= 960.0 + amp + w:
        return 230
    return None
