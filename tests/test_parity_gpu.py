"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the
same seeded inputs.  Bit-exact for voxel indices, occupancy, collision counts,
argmin; scores are expected bit-equal too (same summation tree) and are
additionally checked to 1e-5 relative (north_star tolerance)."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _mk(capi, oracle_mod, k, m, n, d=5, n_noise=1, **kw):
    ctx = capi.Context(k=k, m=m, n=n, d=d, n_noise=n_noise, **kw)
    okw = {a: b for a, b in kw.items() if a not in ("update_mode", "update_gain")}
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, n_noise=n_noise, **okw)
    return ctx, orc


def test_sincos_bit_exact(capi, oracle_mod):
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-8, 8, 20000), rng.uniform(-1e5, 1e5, 2000),
                        [0.0, -0.0, 1e6, -3e7, np.pi / 2, -np.pi / 2]]).astype(np.float32)
    with capi.Context(k=1, m=1, n=1) as ctx:
        s, c = ctx.debug_sincos(x)
    so, co = oracle_mod.sincos(x)
    assert np.array_equal(s.view(np.uint32), so.view(np.uint32))
    assert np.array_equal(c.view(np.uint32), co.view(np.uint32))


@pytest.mark.parametrize("d", [5, 4])
def test_fk_bit_exact(capi, oracle_mod, d):
    rng = np.random.default_rng(1)
    wp = rng.uniform(-3.2, 3.2, (4096, d)).astype(np.float32)
    with capi.Context(k=1, m=1, n=1, d=d) as ctx:
        pts = ctx.debug_fk(wp)
    ref = oracle_mod.fk_points(wp)
    assert np.array_equal(pts.view(np.uint32), ref.view(np.uint32))


def test_noise_philox_bit_exact(capi, oracle_mod):
    ctx, orc = _mk(capi, oracle_mod, 3, 257, 16, sigma=0.1, seed=1234)
    for it in (0, 5):
        a, b = ctx.noise(it), orc.noise(it)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    ctx.close()


@pytest.mark.parametrize("frame", [0, 1])
def test_grid_lookup_indices(capi, oracle_mod, frame):
    g = synth.box_grid(64, 0.02)
    rng = np.random.default_rng(2)
    pts = rng.uniform(-0.7, 1.4, (20000, 3)).astype(np.float32)
    # points on cell boundaries and the frame edges
    edge = np.array([[0.64, 0, 0.5], [-0.64, 0, 0.5], [0, 0.64, 0.5], [0, 0, 1.28], [0, 0, 0],
                     [0.1, -0.1, 0.1], [0.3, 0.1, 0.3], [0.02, 0.04, 0.06]], np.float32)
    pts = np.concatenate([pts, edge])
    ctx, orc = _mk(capi, oracle_mod, 1, 1, 1)
    ctx.set_grid(g, 0.02, frame)
    orc.set_grid(g, 0.02, frame)
    idx, col = ctx.grid_lookup(pts)
    assert np.array_equal(idx, orc.grid_index(pts))
    assert np.array_equal(col, orc.grid_collides(pts))
    assert (idx >= 0).sum() > 1000 and col.sum() > 0
    ctx.close()


CASES = [  # (k, m, n, d, n_noise, grid, sigma)
    (1, 64, 32, 5, 1, "u8", 0.1),      # BASELINE config 1
    (1, 64, 32, 5, 1, "f32", 0.1),
    (3, 37, 50, 4, 1, "u8", 0.3),      # ragged m, d = 4 (reference's shipped d), n = 50
    (2, 33, 33, 5, 33, "u8", 0.2),     # per-waypoint noise, n not a multiple of 32
    (1, 1024, 64, 5, 1, "u8", 0.5),    # BASELINE config 2 shape
    (1, 5, 1, 5, 1, "u8", 0.5),        # single waypoint
    (1, 8, 128, 5, 1, "none", 0.1),    # no grid
]


@pytest.mark.parametrize("k,m,n,d,n_noise,grid,sigma", CASES)
def test_cost_eval_parity(capi, oracle_mod, k, m, n, d, n_noise, grid, sigma):
    ctx, orc = _mk(capi, oracle_mod, k, m, n, d, n_noise, sigma=sigma)
    if grid == "u8":
        g = synth.box_grid(64, 0.02)
    elif grid == "f32":
        g = synth.box_distance_grid(64, 0.02)
    if grid != "none":
        ctx.set_grid(g, 0.02)
        orc.set_grid(g, 0.02)
    start = synth.START[:d]
    base = np.stack([capi.plan_lerp(start, synth.GOALS[c % 2][:d], n) for c in range(k)])
    nz = synth.noise(k, m, n_noise, d, sigma)
    for noise in (nz, None):  # external noise, then in-kernel Philox
        a = ctx.cost_eval(base, start, noise, iteration=3)
        b = orc.cost_eval(base, start, noise, iteration=3)
        assert np.array_equal(a["collisions"], b["collisions"])
        assert a["argmin"] == b["argmin"]
        np.testing.assert_allclose(a["scores"], b["scores"], rtol=RTOL, atol=0)
        assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
        assert a["min_score"] == b["min_score"]
    if grid == "u8":
        assert b["collisions"].sum() > 0  # the obstacle is actually hit
    ctx.close()


def test_as_shipped_mode(capi, oracle_mod):
    ctx, orc = _mk(capi, oracle_mod, 2, 16, 25, 5, collision_mode=0, sigma=1.0)
    g = synth.box_grid(64, 0.02)
    ctx.set_grid(g, 0.02)
    orc.set_grid(g, 0.02)
    base = np.stack([capi.plan_lerp(synth.START, synth.GOALS[c], 25) for c in range(2)])
    nz = synth.noise(2, 16, 1, 5, 1.0)
    a, b = ctx.cost_eval(base, synth.START, nz), orc.cost_eval(base, synth.START, nz)
    assert a["collisions"].sum() == 0 and b["collisions"].sum() == 0
    assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
    ctx.close()


def test_limits_cost(capi, oracle_mod):
    lo, hi = -0.3 * np.ones(5, np.float32), 0.6 * np.ones(5, np.float32)
    ctx, orc = _mk(capi, oracle_mod, 1, 64, 32, 5, limit_weight=2.0, sigma=0.3)
    ctx.set_limits(lo, hi)
    orc.set_limits(lo, hi)
    base = capi.plan_lerp(synth.START, synth.GOALS[0], 32)[None]
    nz = synth.noise(1, 64, 1, 5, 0.3)
    a, b = ctx.cost_eval(base, synth.START, nz), orc.cost_eval(base, synth.START, nz)
    assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
    ctx.close()


def test_against_reference_kernel_golden(capi):
    """CUDA path vs the REFERENCE kernel's own outputs (optimize_trajectories run on a B200 by
    oracle/_ref/ref_harness, tests/golden/ref_outputs.npz): AS_SHIPPED scores from the reference's
    noise, to the float-atomicAdd order of the reference (D17)."""
    import os
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_outputs.npz"))
    f = lambda k: G[k].astype(np.uint32).view(np.float32)
    k, m, n, d = (int(v) for v in G["kernel_shape_kmnd"])
    init = f("kernel_initial_trajectory").reshape(n, d)
    base = np.repeat(init[None], k, 0)
    with capi.Context(k=k, m=m, n=n, d=d, collision_mode=0) as ctx:
        r = ctx.cost_eval(base, f("kernel_start"), f("kernel_noise").reshape(k, m, 1, d))
    ref = f("kernel_scores").reshape(k, m + 1)[:, 1:]
    np.testing.assert_allclose(r["scores"], ref, rtol=2e-6)
    assert r["argmin"] == int(np.argmin(ref))


def test_c1_golden_kat(capi):
    """BASELINE config 1 against the committed known-answer hashes (tests/golden/c1_kat.json)."""
    import hashlib
    import json
    import os
    kat = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "c1_kat.json")))
    with capi.Context(k=1, m=64, n=32, sigma=0.1) as ctx:
        ctx.set_grid(synth.box_grid(64, 0.02), 0.02)
        base = capi.plan_lerp(synth.START, synth.GOALS[0], 32)[None]
        r = ctx.cost_eval(base, synth.START, None, iteration=0)
        nz = ctx.noise(0)
    assert hashlib.sha256(r["scores"].tobytes()).hexdigest() == kat["scores_sha256"]
    assert hashlib.sha256(r["collisions"].tobytes()).hexdigest() == kat["collisions_sha256"]
    assert hashlib.sha256(nz.tobytes()).hexdigest() == kat["noise_sha256"]
    assert r["argmin"] == kat["argmin"] and int(r["collisions"].sum()) == kat["collisions_total"]


PK_CASES = [  # (k, m, n, d, grid, sigma, limits)
    (1, 64, 32, 5, "u8", 0.1, False),
    (2, 37, 50, 4, "u8", 0.3, False),     # ragged m, d = 4
    (1, 1, 16, 5, "u8", 0.1, False),
    (1, 5, 1, 5, "u8", 0.5, False),       # single waypoint: the high half is idle
    (1, 1000, 64, 5, "f32", 0.5, False),
    (3, 129, 33, 5, "u8", 0.3, True),      # odd n: the last step's high half is idle
    (1, 256, 128, 5, "none", 0.2, False),
    # the three geometries of the packed kernel on a 148-SM part: small m above runs 4 parts per
    # sample; these fill the chip with 1 and 2 parts per sample
    (1, 57000, 8, 5, "u8", 0.3, False),
    (1, 30000, 16, 4, "f32", 0.3, True),
    (2, 700, 250, 5, "u8", 0.05, True),    # 4 parts x 64 waypoints (62 in the last), limits
    (1, 40, 6, 5, "u8", 0.3, False),       # 4 parts x 2 waypoints, the last part idle
]


@pytest.mark.parametrize("k,m,n,d,grid,sigma,limits", PK_CASES)
def test_packed_kernel_parity(capi, oracle_mod, monkeypatch, k, m, n, d, grid, sigma, limits):
    """cost_kernel_pk (two waypoints per step in packed fma.rn.f32x2) against the oracle, forced on with
    DOAPP_PK=1 at sizes the oracle finishes in seconds; includes angles that are exactly zero
    (straight-line path valid) and denormal-small (must take the IEEE fallback per half)."""
    monkeypatch.setenv("DOAPP_PK", "1")
    kw = dict(sigma=sigma)
    if limits:
        kw["limit_weight"] = 2.0
    ctx, orc = _mk(capi, oracle_mod, k, m, n, d, 1, **kw)
    if limits:
        lo, hi = -0.3 * np.ones(d, np.float32), 0.6 * np.ones(d, np.float32)
        ctx.set_limits(lo, hi)
        orc.set_limits(lo, hi)
    if grid != "none":
        g = synth.box_grid(64, 0.02) if grid == "u8" else synth.box_distance_grid(64, 0.02)
        ctx.set_grid(g, 0.02)
        orc.set_grid(g, 0.02)
    start = synth.START[:d]
    base = np.stack([capi.plan_lerp(start, synth.GOALS[c % 2][:d], n) for c in range(k)])
    nz = synth.noise(k, m, 1, d, sigma)
    if m >= 8:
        nz[0, 1] = 0.0                      # x == base: zeros for the joints that do not move
        nz[0, 2, 0, :] = 1e-41              # denormal angles: outside the verified ranges
        nz[0, 5, 0, 1] = -base[0, 3, 1]     # one angle exactly 0 in the middle of a trajectory
        nz[0, 6, 0, 0] = 3e5                # |angle| >= 2^17: full-range sin/cos
    for noise in (nz, None):
        a = ctx.cost_eval(base, start, noise, iteration=2)
        b = orc.cost_eval(base, start, noise, iteration=2)
        assert np.array_equal(a["collisions"], b["collisions"])
        assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
        assert a["argmin"] == b["argmin"]
    ctx.close()
