"""Occupancy -> exact Euclidean distance field ("next" row 2 of SURVEY 8f): replaces the host
KD-tree fill CPUDistanceGrid::update (cpu_distance_grid.cu:87-202); the consumer is the
reference's sphere test `distance <= COLLISION_DISTANCE` (generate_trajectories.cu:389,412)."""
import numpy as np
import pytest

from tests import synth


def test_oracle_scan_equals_brute_force(oracle_mod):
    rng = np.random.default_rng(3)
    for shape, p in (((9, 11, 13), 0.02), ((12, 12, 12), 0.2), ((6, 5, 4), 0.0), ((10, 7, 5), 0.01),
                     ((8, 8, 8), 1.0), ((1, 1, 40), 0.1)):
        orc = oracle_mod.Oracle(k=1, m=1, n=1)
        orc.set_grid((rng.random(shape) < p).astype(np.uint8), 0.02)
        a, b = orc.distance_field(), orc.distance_field(brute=True)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    # the synthetic box: distance 0 inside, one voxel step = one resolution next to a face
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    g = synth.box_grid(64, 0.02)
    orc.set_grid(g, 0.02)
    d = orc.distance_field()
    assert np.all(d[g != 0] == 0) and np.all(d[g == 0] > 0)
    iz, iy, ix = np.argwhere(g != 0)[0]
    assert d[iz, iy, ix - 1] == np.float32(0.02)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,res,fill", [((64, 64, 64), 0.02, "box"), ((128, 128, 128), 0.01, "box"),
                                            ((40, 50, 70), 0.02, 0.001), ((33, 65, 97), 0.01, 0.05),
                                            ((32, 32, 32), 0.02, 0.0), ((16, 16, 300), 0.01, 0.002),
                                            # lines of 129..256 voxels (eight outputs per lane of the
                                            # exhaustive kernel), and above 256 (the scan kernel)
                                            ((200, 40, 33), 0.01, 0.003), ((256, 130, 20), 0.01, 0.0005),
                                            ((300, 20, 21), 0.01, 0.004), ((20, 257, 31), 0.01, 0.004),
                                            ((128, 128, 128), 0.01, "cloud")])
def test_distance_field_parity(capi, oracle_mod, shape, res, fill):
    rng = np.random.default_rng(7)
    if fill == "box":
        g = synth.box_grid(shape[0], res)
    elif fill == "cloud":  # BASELINE config 3's grid: a 300k-point cloud scattered into 128^3
        o = oracle_mod.Oracle(k=1, m=1, n=1)
        o.set_grid(np.zeros(shape, np.uint8), res)
        o.scatter(synth.cloud(300_000, 0, size=shape[0] * res), 1, True)
        g = o.cells.copy()
    else:
        g = (rng.random(shape) < fill).astype(np.uint8)
    ctx = capi.Context(k=1, m=1, n=1)
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    ctx.set_grid(g, res)
    orc.set_grid(g, res)
    ctx.grid_distance_field()
    got = ctx.grid_download(shape, np.float32)
    assert np.array_equal(got.view(np.uint32), orc.distance_field().view(np.uint32))
    ctx.close()


@pytest.mark.gpu
def test_cost_on_distance_field(capi, oracle_mod):
    """The cost path on the derived F32 grid (collision iff distance <= 0.05, gt.cu:412) equals
    the oracle on the oracle's distance field, and sees MORE collisions than the occupancy test."""
    k, m, n = 1, 256, 32
    g = synth.box_grid(64, 0.02)
    ctx = capi.Context(k=k, m=m, n=n, sigma=0.1)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, sigma=0.1)
    ctx.set_grid(g, 0.02)
    orc.set_grid(g, 0.02)
    base = capi.plan_lerp(synth.START, synth.GOALS[0], n)[None]
    occ = ctx.cost_eval(base, synth.START, None, iteration=1)
    ctx.grid_distance_field()
    orc.set_grid(orc.distance_field(), 0.02)
    a = ctx.cost_eval(base, synth.START, None, iteration=1)
    b = orc.cost_eval(base, synth.START, None, iteration=1)
    assert np.array_equal(a["collisions"], b["collisions"])
    assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert a["collisions"].sum() > occ["collisions"].sum() > 0
    ctx.close()
