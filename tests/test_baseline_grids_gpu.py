"""GPU parity on the BASELINE grids themselves (VERDICT r01 "weak #1"): the fast voxel-index path
(collides<..., FIDX>: per-axis bit-pattern thresholds found by bisection at doapp_grid_alloc and a
reciprocal divide verified per resolution) is resolution-specific code, so it is compared with the
oracle at res 0.01 (C2 / C4: 128^3) and 0.0025 (C5: 512^3, 5 % random fill), not only at the
64^3 / 0.02 grid of the other parity tests.  Lookup spec: cpu_distance_grid.cu:113-126."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def _pair(capi, oracle_mod, k, m, n, **kw):
    return (capi.Context(k=k, m=m, n=n, d=5, **kw), oracle_mod.Oracle(k=k, m=m, n=n, d=5, **kw))


def _same_eval(a, b):
    assert np.array_equal(a["collisions"], b["collisions"])
    assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert a["argmin"] == b["argmin"]
    assert a["min_score"] == b["min_score"]


def _same_plan(ctx, a, b):
    st = ctx.plan_state()
    assert np.array_equal(st["found_better"], b["found_better"])
    assert np.array_equal(st["trajectories"].view(np.uint32), b["trajectories"].view(np.uint32))
    assert np.array_equal(st["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert np.float32(a["best_score"]) == np.float32(b["best_score"])
    assert np.array_equal(a["best_traj"].view(np.uint32), b["best_traj"].view(np.uint32))
    assert np.array_equal(a["best_vel"].view(np.uint32), b["best_vel"].view(np.uint32))


@pytest.mark.parametrize("kind", ["u8", "f32"])
def test_c2_grid_cost_eval_and_100_iterations(capi, oracle_mod, kind):
    """BASELINE config 2 as benchmarked: 128^3 @ 1 cm, 64 waypoints x 1024 samples, 100
    iterations of the persistent planner."""
    g = synth.box_grid(128, 0.01) if kind == "u8" else synth.box_distance_grid(128, 0.01)
    ctx, orc = _pair(capi, oracle_mod, 1, 1024, 64, sigma=0.05, select_mode=1)
    ctx.set_grid(g, 0.01)
    orc.set_grid(g, 0.01)
    start, goal = synth.START, synth.GOALS[1]
    init = capi.plan_lerp(start, goal, 64)
    a = ctx.cost_eval(init[None], start, None, iteration=7)
    b = orc.cost_eval(init[None], start, None, iteration=7)
    _same_eval(a, b)
    assert b["collisions"].sum() > 0
    pa = ctx.plan(start, goal, init, 100)
    assert ctx.plan_mode() & 1, "config 2 must take the persistent planner"
    pb = orc.plan(start, init, 100, threads=8)
    _same_plan(ctx, pa, pb)
    assert pb["found_better"][0] == 1
    ctx.close()


def test_c2_sigma_large_many_updates(capi, oracle_mod):
    """Many accepted updates (every mis-speculated pass of the pipelined persistent loop is
    re-evaluated): a start far from the optimum, larger sigma, 2 climbers, 60 iterations."""
    g = synth.box_grid(128, 0.01)
    ctx, orc = _pair(capi, oracle_mod, 2, 1024, 64, sigma=0.2, select_mode=1)
    ctx.set_grid(g, 0.01)
    orc.set_grid(g, 0.01)
    start, goal = synth.START, synth.GOALS[1]
    rng = np.random.default_rng(3)
    init = (capi.plan_lerp(start, goal, 64) + 0.3 * rng.standard_normal((64, 5))).astype(np.float32)
    pa = ctx.plan(start, goal, init, 60)
    pb = orc.plan(start, init, 60, threads=8)
    _same_plan(ctx, pa, pb)
    ctx.close()


def test_c4_slice_inside_1m_run(capi, oracle_mod):
    """BASELINE config 4: n = 128, packed one-thread-per-sample kernel (16 climbers x 4096 samples
    fill the chip, so T = 1), the samples a slice at an odd offset inside a 1M-sample run, Philox
    noise keyed by the GLOBAL sample index."""
    g = synth.box_grid(128, 0.01)
    kw = dict(sigma=0.05, select_mode=1, sample_offset=777_001,
              global_samples_per_climber=1_048_576)
    ctx, orc = _pair(capi, oracle_mod, 16, 4096, 128, **kw)
    ctx.set_grid(g, 0.01)
    orc.set_grid(g, 0.01)
    start = synth.START
    base = np.stack([capi.plan_lerp(start, synth.GOALS[c % 2], 128) for c in range(16)])
    a = ctx.cost_eval(base, start, None, iteration=2)
    b = orc.cost_eval(base, start, None, iteration=2, threads=8)
    _same_eval(a, b)
    assert b["collisions"].sum() > 0
    ctx.close()


def test_c4_plan_sharded_slices_merge(capi, oracle_mod):
    """Two shards of one 8192-sample run planned separately for one iteration: the better of the
    two shard winners is the winner of the whole run (host argmin planning_state.cu:113-122)."""
    g = synth.box_grid(128, 0.01)
    start, goal = synth.START, synth.GOALS[1]
    init = capi.plan_lerp(start, goal, 128)
    whole = oracle_mod.Oracle(k=1, m=8192, n=128, d=5, sigma=0.05, select_mode=1)
    whole.set_grid(g, 0.01)
    w = whole.cost_eval(init[None], start, None, iteration=0, threads=8)
    best = []
    for off in (0, 4096):
        with capi.Context(k=1, m=4096, n=128, d=5, sigma=0.05, select_mode=1, sample_offset=off,
                          global_samples_per_climber=8192) as ctx:
            ctx.set_grid(g, 0.01)
            a = ctx.cost_eval(init[None], start, None, iteration=0)
            assert np.array_equal(a["scores"].view(np.uint32),
                                  w["scores"][:, off:off + 4096].view(np.uint32))
            best.append((a["min_score"], a["argmin"]))
    assert min(best) == (w["min_score"], w["argmin"])


@pytest.mark.parametrize("pk", ["1", "0"])
def test_c5_grid_slice(capi, oracle_mod, monkeypatch, pk):
    """BASELINE config 5: 512^3 @ 2.5 mm with 5 % random fill, sigma 0.5, a 2048-sample slice;
    DOAPP_PK=1 takes cost_kernel_pk<..., MP> with 4 parts per sample (the geometry C5 runs with
    at 16384 samples), DOAPP_PK=0 the clustered scalar kernel."""
    monkeypatch.setenv("DOAPP_PK", pk)
    g = synth.box_grid(512, 0.0025, fill=0.05)
    kw = dict(sigma=0.5, select_mode=1, sample_offset=5000, global_samples_per_climber=16384)
    ctx, orc = _pair(capi, oracle_mod, 1, 2048, 128, **kw)
    ctx.set_grid(g, 0.0025)
    orc.set_grid(g, 0.0025)
    start, goal = synth.START, synth.GOALS[1]
    init = capi.plan_lerp(start, goal, 128)
    a = ctx.cost_eval(init[None], start, None, iteration=1)
    b = orc.cost_eval(init[None], start, None, iteration=1, threads=8)
    _same_eval(a, b)
    assert b["collisions"].sum() > 100  # the random fill is hit
    pa = ctx.plan(start, goal, init, 3)
    pb = orc.plan(start, init, 3, threads=8)
    _same_plan(ctx, pa, pb)
    ctx.close()


def test_c5_lookup_indices(capi, oracle_mod):
    """Voxel indices at res 0.0025 on points at and around cell boundaries and the frame edges."""
    n, res = 512, 0.0025
    g = np.zeros((n, n, n), np.uint8)
    rng = np.random.default_rng(11)
    pts = rng.uniform(-0.7, 1.35, (20000, 3)).astype(np.float32)
    k = rng.integers(-260, 520, (6000, 3))
    onb = (k * np.float32(res)).astype(np.float32)            # multiples of the resolution
    nxt = np.nextafter(onb, np.float32(np.inf)).astype(np.float32)
    prv = np.nextafter(onb, np.float32(-np.inf)).astype(np.float32)
    pts = np.concatenate([pts, onb, nxt, prv])
    with capi.Context(k=1, m=1, n=1) as ctx:
        orc = oracle_mod.Oracle(k=1, m=1, n=1)
        ctx.set_grid(g, res)
        orc.set_grid(g, res)
        idx, _ = ctx.grid_lookup(pts)
        assert np.array_equal(idx, orc.grid_index(pts))
        assert (idx >= 0).sum() > 5000


def test_replan_after_grid_realloc(capi, oracle_mod):
    """the replan CUDA-graph cache must not replay a graph captured for another grid.
    Re-allocating a smaller grid (same buffer, new dims / resolution) between two
    doapp_replan_dynamic calls must give the result of a fresh context."""
    k, m, n, iters, P = 1, 128, 32, 3, 20000
    init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
    clouds = np.stack([synth.cloud(P, it) for it in range(iters)])
    with capi.Context(k=k, m=m, n=n, sigma=0.05, select_mode=1) as ctx:
        ctx.set_grid(np.zeros((128, 128, 128), np.uint8), 0.01)
        ctx.replan_dynamic(synth.START, synth.GOALS[1], init, clouds)
        ctx.set_grid(np.zeros((64, 64, 64), np.uint8), 0.02)  # fits the old capacity
        r2 = ctx.replan_dynamic(synth.START, synth.GOALS[1], init, clouds)
        got = ctx.grid_download((64, 64, 64))
    with capi.Context(k=k, m=m, n=n, sigma=0.05, select_mode=1) as fresh:
        fresh.set_grid(np.zeros((64, 64, 64), np.uint8), 0.02)
        r3 = fresh.replan_dynamic(synth.START, synth.GOALS[1], init, clouds)
        want = fresh.grid_download((64, 64, 64))
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    orc.set_grid(np.zeros((64, 64, 64), np.uint8), 0.02)
    orc.scatter(clouds[-1], 1, True)
    assert np.array_equal(got, want) and np.array_equal(got, orc.cells)
    assert np.array_equal(r2["best_traj"].view(np.uint32), r3["best_traj"].view(np.uint32))
    assert np.float32(r2["best_score"]) == np.float32(r3["best_score"])
