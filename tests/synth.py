"""Seeded synthetic inputs of the BASELINE configs, shared by
tests, smoke() and bench.py.  numpy only."""
import numpy as np

START = np.zeros(5, np.float32)
GOALS = [np.array([0, 1.0, 0, 0.5, 0], np.float32),       # task/src/main.cpp:35 + wrist_rotate 0
         np.array([0, -1.0, 1.5, -1.5, 0], np.float32)]   # task/src/main.cpp:39


def box_grid(n: int, res: float, lo=(0.10, -0.10, 0.10), hi=(0.30, 0.10, 0.30),
             fill: float = 0.0, seed: int = 7) -> np.ndarray:
    """uint8 occupancy [iz, iy, ix] in the CDG frame (x,y centred, z from 0) with one
    axis-aligned box obstacle; `fill` adds that fraction of random occupied voxels."""
    g = np.zeros((n, n, n), np.uint8)
    c = (np.arange(n) + 0.5) * res
    x = c - 0.5 * n * res
    y = c - 0.5 * n * res
    z = c
    mx = (x >= lo[0]) & (x < hi[0])
    my = (y >= lo[1]) & (y < hi[1])
    mz = (z >= lo[2]) & (z < hi[2])
    g[np.ix_(mz, my, mx)] = 1
    if fill > 0:
        rng = np.random.Generator(np.random.Philox(seed))
        for iz in range(n):  # slice-wise keeps peak memory small at 512^3
            g[iz] |= (rng.random((n, n), np.float32) < fill).astype(np.uint8)
    return g


def box_distance_grid(n: int, res: float, lo=(0.10, -0.10, 0.10), hi=(0.30, 0.10, 0.30)):
    """fp32 Euclidean distance to the box, same frame/dims as box_grid."""
    c = (np.arange(n) + 0.5) * res
    x = (c - 0.5 * n * res)[None, None, :]
    y = (c - 0.5 * n * res)[None, :, None]
    z = c[:, None, None]
    dx = np.maximum(np.maximum(lo[0] - x, x - hi[0]), 0)
    dy = np.maximum(np.maximum(lo[1] - y, y - hi[1]), 0)
    dz = np.maximum(np.maximum(lo[2] - z, z - hi[2]), 0)
    return np.sqrt(dx * dx + dy * dy + dz * dz).astype(np.float32)


def noise(k, m, n_noise, d, sigma, seed=1234):
    rng = np.random.Generator(np.random.Philox(seed))
    return (sigma * rng.standard_normal((k, m, n_noise, d))).astype(np.float32)


def cloud(num_points: int, iteration: int, size: float = 1.28, seed: int = 99):
    """Surface samples of a box translating 1 cm per iteration + 10 % outliers (§8d)."""
    rng = np.random.Generator(np.random.Philox(key=seed + 1000003 * iteration))
    nb = int(num_points * 0.9)
    lo = np.array([0.10 + 0.01 * iteration, -0.10, 0.10], np.float32)
    ext = np.array([0.20, 0.20, 0.20], np.float32)
    p = rng.random((nb, 3), np.float32)
    face = rng.integers(0, 6, nb)
    ax = face % 3
    p[np.arange(nb), ax] = (face // 3).astype(np.float32)
    pts = lo + p * ext
    out = rng.random((num_points - nb, 3), np.float32) * size
    out[:, :2] -= size / 2
    return np.concatenate([pts, out]).astype(np.float32)
