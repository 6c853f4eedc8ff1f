"""Depth frames -> occupancy grid ("next" row 1 of SURVEY 8f): mapping_node.cpp:89-137
(deprojection), :195 (transform), :262-263 (crop), og.cpp:47-58 (set), fused in one kernel."""
import numpy as np
import pytest

from tests import synth


def _numpy_restatement(depth, intr, extr, crop):
    """The same arithmetic in numpy float32, operation by operation (mapping_node.cpp:110-135)."""
    f = np.float32
    pts = []
    for c in range(depth.shape[0]):
        h, w = depth[c].shape
        v, u = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        d = depth[c]
        ok = d != 0
        df = d.astype(f)
        cx, cy = f(intr[c, 2]), f(intr[c, 3])
        kx = f(np.float64(f(0.001)) / intr[c, 0])
        ky = f(np.float64(f(0.001)) / intr[c, 1])
        x = ((u.astype(f) - cx) * df) * kx
        y = ((v.astype(f) - cy) * df) * ky
        z = df * f(0.001)
        r, t = extr[c, :9].reshape(3, 3), extr[c, 9:]
        q = [((r[a, 0] * x + r[a, 1] * y) + r[a, 2] * z) + t[a] for a in range(3)]
        half, full = f(crop / 2.0), f(crop)
        keep = ok & (q[0] >= -half) & (q[0] <= half) & (q[1] >= -half) & (q[1] <= half) & \
            (q[2] >= 0) & (q[2] <= full)
        pts.append(np.stack([q[0][keep], q[1][keep], q[2][keep]], axis=1))
    return np.concatenate(pts).astype(f)


def test_oracle_depth_scatter_matches_numpy_and_point_scatter(oracle_mod):
    depth, intr, extr = synth.depth_rig(3, 160, 90)
    n, res = 64, 0.02
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    orc.set_grid(np.zeros((n, n, n), np.uint8), res)
    kept = orc.depth_scatter(depth, intr, extr, n * res)
    got = orc.cells.copy()
    pts = _numpy_restatement(depth, intr, extr, n * res)
    assert kept == len(pts) and kept > 1000
    orc.set_grid(np.zeros((n, n, n), np.uint8), res)
    orc.scatter(pts, 1, True)
    assert np.array_equal(got, orc.cells)
    assert got.sum() > 200
    # an empty frame set leaves the grid clear; accumulation without clearing keeps old voxels
    assert orc.depth_scatter(np.zeros_like(depth), intr, extr, n * res) == 0 and orc.cells.sum() == 0


@pytest.mark.gpu
@pytest.mark.parametrize("cams,w,h,n,res,frame", [(3, 480, 270, 128, 0.01, 0), (1, 64, 48, 64, 0.02, 0),
                                                  (2, 160, 90, 64, 0.02, 1)])
def test_depth_to_grid_parity(capi, oracle_mod, cams, w, h, n, res, frame):
    depth, intr, extr = synth.depth_rig(cams, w, h)
    if frame == 1:  # the OccupancyGrid cube is centred on z too (og.cpp:29-40)
        extr[:, 11] -= 0.0
    ctx = capi.Context(k=1, m=1, n=1)
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    z = np.zeros((n, n, n), np.uint8)
    ctx.set_grid(z, res, frame)
    orc.set_grid(z.copy(), res, frame)
    ka = ctx.grid_from_depth(depth, intr, extr, n * res)
    kb = orc.depth_scatter(depth, intr, extr, n * res)
    got = ctx.grid_download((n, n, n))
    assert ka == kb and ka > 100
    assert np.array_equal(got, orc.cells)
    assert got.sum() > 50
    # accumulate a second, shifted view without clearing; then all-invalid frames + clear
    d2 = np.roll(depth, 7, axis=2)
    assert ctx.grid_from_depth(d2, intr, extr, n * res, val=3, clear_first=False) == \
        orc.depth_scatter(d2, intr, extr, n * res, val=3, clear_first=False)
    assert np.array_equal(ctx.grid_download((n, n, n)), orc.cells)
    assert ctx.grid_from_depth(np.zeros_like(depth), intr, extr, n * res) == 0
    assert ctx.grid_download((n, n, n)).sum() == 0
    ctx.close()
