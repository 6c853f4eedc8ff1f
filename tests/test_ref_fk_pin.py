"""Stage 2 (forward kinematics + collision count) pinned against the REFERENCE's own device code.

tests/golden/ref_fk_outputs.npz holds outputs of generate_trajectories.cu:303-387 (init_rotx/y/z,
init_rot_rpy, eye) and :391-466 (num_collisions, with the single stub line :394 dropped), compiled
unmodified into oracle/_ref/ref_fk_harness and run on a B200 (oracle/ref_fk_harness.cu,
tests/golden/make_ref_fk_golden.py).  The oracle's restatement must reproduce them: the matrix
STRUCTURE exactly (which entries are 0, 1, +-sin, cos: axes and signs), the values to <= 2 ulp
(CUDA's sinf/cosf vs the deterministic sin/cos both sides of this repo share), and the collision
counts of the as-shipped chain (Matrix*Vector = row sums) exactly wherever no link
point sits within a hair of a cell boundary.  That pins the axes order (y, z, y, y, y), the
rpy composition, the link order, the `<= 0.05f` predicate and the grid frame."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def ref():
    z = np.load(os.path.join(HERE, "golden", "ref_fk_outputs.npz"))
    return {k: z[k] for k in z.files}


def _f(a):
    return np.asarray(a, np.uint32).view(np.float32)


def _ulps(a, b):
    a = np.asarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.asarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = np.where(b < 0, -(b & 0x7FFFFFFF), b)
    return np.abs(a - b)


def test_rotation_builders_match_reference_device_code(oracle_mod, ref):
    ang = _f(ref["rot_angles"])
    for axis, key in enumerate(("rot_x", "rot_y", "rot_z")):
        got = _f(ref[key]).reshape(-1, 3, 3)
        want = np.stack([oracle_mod.rotation(axis, a) for a in ang])
        # structure: exact zeros and ones where gt.cu:303-352 writes literals
        lit = np.zeros((3, 3), bool)
        lit[axis, :] = lit[:, axis] = True
        assert np.array_equal(got[:, lit].view(np.uint32), want[:, lit].view(np.uint32))
        # sin / cos entries: |value| <= 1, so an absolute 2.4e-7 is 2 ulp at 1.0
        np.testing.assert_allclose(got[:, ~lit], want[:, ~lit], rtol=0, atol=2.4e-7)
        # signs of the sine entries (the sense of rotation) away from sin ~ 0
        big = np.abs(np.sin(ang.astype(np.float64))) > 1e-3
        assert np.array_equal(np.sign(got[big][:, ~lit]), np.sign(want[big][:, ~lit]))


def test_rpy_and_eye_match_reference_device_code(oracle_mod, ref):
    got = _f(ref["rot_rpy"]).reshape(3, 3)
    want = oracle_mod.rotation_rpy(0.0, -1.5707963267948966, 0.0)  # the call of gt.cu:456
    np.testing.assert_allclose(got, want, rtol=0, atol=1.2e-7)
    assert np.array_equal(np.round(got), np.round(want))  # [[0,0,-1],[0,1,0],[1,0,0]]
    assert np.array_equal(_f(ref["rot_eye"]).reshape(3, 3), np.eye(3, dtype=np.float32))


def _hash32(x):
    x = x.astype(np.uint64)
    M = np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & M
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & M
    x ^= x >> np.uint64(16)
    return x


def test_num_collisions_matches_reference_device_code(oracle_mod, ref):
    n, seed = (int(v) for v in ref["coll_grid_n_seed"])
    res = float(_f(ref["coll_grid_res"])[0])
    wps = _f(ref["coll_waypoints"]).reshape(-1, 5)
    counts = ref["coll_counts"]
    # the planted grid of oracle/ref_fk_harness.cu: cell value by a hash of its logical index
    idx = np.arange(n ** 3, dtype=np.uint64)
    h = _hash32(idx ^ np.uint64(seed)) % np.uint64(6)
    vals = np.full(n ** 3, 1.0, np.float32)
    vals[h == 0] = 0.0
    vals[h == 1] = np.float32(0.05)
    vals[h == 2] = np.float32(0.050000004)
    grid = vals.reshape(n, n, n)
    orc = oracle_mod.Oracle(k=1, m=1, n=1, d=5)  # collision_distance 0.05 (gt.cu:389)
    orc.set_grid(grid, res)
    pts = oracle_mod.fk_points(wps, as_shipped_matvec=True).reshape(-1, 5, 3)
    col = orc.grid_collides(pts.reshape(-1, 3)).reshape(-1, 5)
    mine = col.sum(axis=1)
    # a point within 1e-5 cells of a cell boundary may land either side: CUDA's sinf/cosf and the
    # deterministic sin/cos differ in the last ulp
    cell = pts.astype(np.float64) / res
    frac = np.abs(cell - np.round(cell))
    # coordinates that do not depend on an angle are the same numbers on both sides (as shipped
    # the row sums give y == 1 after a roty, z == 1 after the rotz, and a constant last point)
    const = np.zeros((5, 3), bool)
    const[[0, 2, 3], 1] = True
    const[1, 2] = True
    const[4, :] = True
    safe = ((frac > 1e-5) | const[None]).all(axis=(1, 2))
    assert safe.mean() > 0.99
    assert np.array_equal(mine[safe], counts[safe])
    assert (mine == counts).mean() > 0.999
    assert len(np.unique(counts)) >= 4 and counts[0] == mine[0]  # varied counts; the zero pose
