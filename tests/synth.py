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


def depth_rig(cams: int = 3, w: int = 480, h: int = 270, seed: int = 5):
    """Synthetic RealSense-style rig (mapping/launch/cameras.launch:13-14: 480x270 depth):
    uint16 millimetre depth frames [cams][h][w] (a far wall, a nearer box, 5 % invalid zeros),
    pinhole intrinsics [cams][4] = fx, fy, cx, cy and camera -> arm extrinsics [cams][12]
    (row-major rotation, translation) of cameras on a ring looking at the workspace centre."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    depth = np.empty((cams, h, w), np.uint16)
    intr = np.empty((cams, 4), np.float64)
    extr = np.empty((cams, 12), np.float32)
    target = np.array([0.0, 0.0, 0.3])
    for c in range(cams):
        d = 700.0 + 40.0 * rng.standard_normal((h, w))
        v0, u0 = h // 4 + 10 * c, w // 3 + 15 * c
        d[v0:v0 + h // 3, u0:u0 + w // 4] = 450.0 + 5.0 * rng.standard_normal((h // 3, w // 4))
        d = np.clip(d, 1, 65535)
        d[rng.random((h, w)) < 0.05] = 0
        depth[c] = d.astype(np.uint16)
        intr[c] = (240.0 + 3.0 * c, 241.5 + 2.0 * c, w / 2 - 0.5 + 0.25 * c, h / 2 - 0.5 - 0.25 * c)
        ang = 2.0 * np.pi * c / cams + 0.3
        pos = np.array([0.6 * np.cos(ang), 0.6 * np.sin(ang), 0.55])
        zc = target - pos
        zc /= np.linalg.norm(zc)                       # optical axis
        xc = np.cross(zc, np.array([0.0, 0.0, 1.0]))
        xc /= np.linalg.norm(xc)
        yc = np.cross(zc, xc)
        rot = np.stack([xc, yc, zc], axis=1)           # columns: camera axes in the arm frame
        extr[c, :9] = rot.astype(np.float32).ravel()
        extr[c, 9:] = pos.astype(np.float32)
    return depth, intr, extr
