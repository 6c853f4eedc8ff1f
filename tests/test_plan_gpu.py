"""GPU parity of the planner loop (keep-best, gt.cu:144-258 with D5/D6 fixed),
the voxel scatter (og.cpp:42-58) and the dynamic replan, against the oracle."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def _pair(capi, oracle_mod, k, m, n, d=5, n_noise=1, **kw):
    return (capi.Context(k=k, m=m, n=n, d=d, n_noise=n_noise, **kw),
            oracle_mod.Oracle(k=k, m=m, n=n, d=d, n_noise=n_noise, **kw))


@pytest.mark.parametrize("k,m,n,d,n_noise,select,ext", [
    (1, 64, 32, 5, 1, 0, True),
    (4, 48, 32, 5, 1, 1, True),
    (2, 100, 50, 4, 1, 0, False),
    (2, 40, 33, 5, 33, 1, False),
    (1, 1024, 64, 5, 1, 1, False),
])
def test_plan_parity(capi, oracle_mod, k, m, n, d, n_noise, select, ext):
    iters = 6
    ctx, orc = _pair(capi, oracle_mod, k, m, n, d, n_noise, sigma=0.05, select_mode=select)
    g = synth.box_grid(64, 0.02)
    ctx.set_grid(g, 0.02)
    orc.set_grid(g, 0.02)
    start, goal = synth.START[:d], synth.GOALS[0][:d]
    init = capi.plan_lerp(start, goal, n)
    nz = synth.noise(iters * k, m, n_noise, d, 0.05).reshape(iters, k, m, n_noise, d) if ext else None
    a = ctx.plan(start, goal, init, iters, nz)
    b = orc.plan(start, init, iters, nz)
    st = ctx.plan_state()
    assert np.array_equal(st["found_better"], b["found_better"])
    assert np.array_equal(st["trajectories"].view(np.uint32), b["trajectories"].view(np.uint32))
    assert np.array_equal(st["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert np.array_equal(st["rounded"], b["rounded"])
    assert a["best_climber"] == b["best_climber"]
    assert np.float32(a["best_score"]) == np.float32(b["best_score"])
    assert np.array_equal(a["best_traj"].view(np.uint32), b["best_traj"].view(np.uint32))
    assert np.array_equal(a["best_vel"].view(np.uint32), b["best_vel"].view(np.uint32))
    # idempotence: planning again from the same inputs gives the same answer
    a2 = ctx.plan(start, goal, init, iters, nz)
    assert np.array_equal(a2["best_traj"], a["best_traj"])
    ctx.close()


def test_plan_improves(capi, oracle_mod):
    ctx, orc = _pair(capi, oracle_mod, 2, 512, 64, sigma=0.05, select_mode=1)
    start, goal = synth.START, synth.GOALS[1]
    init = capi.plan_lerp(start, goal, 64)
    g = synth.box_grid(128, 0.01)
    ctx.set_grid(g, 0.01)
    orc.set_grid(g, 0.01)
    s0 = orc.cost_sample(init, start, np.zeros((1, 5), np.float32))[0]
    a = ctx.plan(start, goal, init, 20)
    assert a["best_score"] <= s0
    ctx.close()


@pytest.mark.parametrize("frame,n,res", [(1, 64, 0.02), (0, 128, 0.01)])
def test_scatter_parity(capi, oracle_mod, frame, n, res):
    ctx, orc = _pair(capi, oracle_mod, 1, 1, 1)
    pts = synth.cloud(300_000, 0, size=n * res)
    if frame == 1:  # OG frame is centred on z too
        pts[:, 2] -= 0.5 * n * res
    # points exactly on / just outside the frame edges
    pts[:8] = np.array([[n * res / 2, 0, 0.1], [-n * res / 2, 0, 0.1], [0, 0, n * res], [0, 0, -1e-9],
                        [1e9, 0, 0], [np.nan, 0, 0], [0, np.inf, 0], [0, 0, 0]], np.float32)
    z = np.zeros((n, n, n), np.uint8)
    ctx.set_grid(z, res, frame)
    orc.set_grid(z.copy(), res, frame)
    ctx.scatter(pts, 1, True)
    orc.scatter(pts, 1, True)
    got = ctx.grid_download((n, n, n))
    assert np.array_equal(got, orc.cells)
    assert got.sum() > 1000
    # accumulate without clearing, different value; then clear + empty cloud
    p2 = synth.cloud(5000, 7, size=n * res)
    ctx.scatter(p2, 3, False)
    orc.scatter(p2, 3, False)
    assert np.array_equal(ctx.grid_download((n, n, n)), orc.cells)
    ctx.scatter(np.zeros((0, 3), np.float32), 1, True)
    assert ctx.grid_download((n, n, n)).sum() == 0
    orc.scatter(np.zeros((0, 3), np.float32), 1, True)
    for cnt in (1, 3, 5, 4099):  # tails of the four-points-per-thread loads
        p3 = synth.cloud(cnt + 10, 11 + cnt, size=n * res)[:cnt]
        ctx.scatter(p3, 2, False)
        orc.scatter(p3, 2, False)
        assert np.array_equal(ctx.grid_download((n, n, n)), orc.cells)
    ctx.close()


def test_grid_from_msg(capi, oracle_mod):
    """OccupancyGrid.msg fields (size, granularity, int8[] grid) -> device grid -> lookups."""
    n, gran = 32, 0.04
    size = n * gran
    rng = np.random.default_rng(5)
    cells = (rng.random((n, n, n)) < 0.1).astype(np.int8)
    ctx, orc = _pair(capi, oracle_mod, 1, 1, 1)
    ctx.grid_from_msg(size, gran, cells.ravel())
    orc.set_grid(cells, gran, 1)
    pts = rng.uniform(-0.7, 0.7, (5000, 3)).astype(np.float32)
    idx, col = ctx.grid_lookup(pts)
    assert np.array_equal(idx, orc.grid_index(pts))
    assert np.array_equal(col, orc.grid_collides(pts))
    assert np.array_equal(ctx.grid_download((n, n, n), np.int8), cells)
    ctx.close()


def test_replan_dynamic_matches_stepwise(capi, oracle_mod):
    """config 3 at reduced size: scatter + one planner iteration per cloud; the final grid must
    equal the oracle's scatter of the last cloud and the score must match an oracle re-evaluation."""
    k, m, n, iters, P = 1, 128, 32, 4, 20000
    ctx, orc = _pair(capi, oracle_mod, k, m, n, sigma=0.05, select_mode=1)
    z = np.zeros((64, 64, 64), np.uint8)
    ctx.set_grid(z, 0.02)
    orc.set_grid(z.copy(), 0.02)
    clouds = np.stack([synth.cloud(P, it) for it in range(iters)])
    init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
    r = ctx.replan_dynamic(synth.START, synth.GOALS[1], init, clouds)
    orc.scatter(clouds[-1], 1, True)
    assert np.array_equal(ctx.grid_download((64, 64, 64)), orc.cells)
    s, _ = orc.cost_sample(r["best_traj"], synth.START, np.zeros((1, 5), np.float32))
    assert np.float32(s) == np.float32(r["best_score"])
    ctx.close()


@pytest.mark.parametrize("env,persist", [("DOAPP_NO_GRAPH", "0"), ("DOAPP_EXACT_DIV", "0"),
                                         ("DOAPP_EXACT_DIV", "1"), ("DOAPP_PK", "0"),
                                         ("DOAPP_PERSIST", None)])
def test_alternate_paths_bit_identical(capi, monkeypatch, env, persist):
    """The un-graphed launch path, the IEEE-divide kernels and the packed two-waypoints-per-step
    kernel (fma.rn.f32x2) give the same bits as the default path; so do the persistent
    one-launch planner (DOAPP_PERSIST=1) and the launch-per-iteration graph (=0)."""
    k, m, n, iters = 3, 200, 64, 6
    init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
    g = synth.box_grid(64, 0.02)
    out = []
    if persist is not None:
        monkeypatch.setenv("DOAPP_PERSIST", persist)
    for flag in ("0", "1"):
        monkeypatch.setenv(env, flag)
        with capi.Context(k=k, m=m, n=n, sigma=0.05) as ctx:
            ctx.set_grid(g, 0.02)
            r = ctx.plan(synth.START, synth.GOALS[1], init, iters)
            st = ctx.plan_state()
            out.append((r["best_traj"].copy(), st["trajectories"].copy(), st["scores"].copy(),
                        st["found_better"].copy(), ctx.launch_count()))
    for a, b in zip(out[0][:4], out[1][:4]):
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8))
