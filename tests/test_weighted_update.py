"""Weighted update rule: compute_update_vector / update_trajectories,
generate_trajectories.cu:117-142 (dead code in the reference, call site :242-251), pinned by the
reference's host loop generate_trajectories.test.cu:200-208.

The reference sums the terms with unordered float atomicAdd; oracle and CUDA kernel both round
every term to a multiple of 2^-28 and sum 64-bit integers, which makes the result independent of
order, launch geometry and GPU count."""
import numpy as np
import pytest

from tests import synth


def _host_loop_update(scores, noise, best):
    """generate_trajectories.test.cu:202-208 for one climber, float accumulation in sample order."""
    m, d = noise.shape
    upd = np.zeros(d, np.float32)
    for t in range(m):
        w = np.float32(np.float32(best) / np.ceil(scores[t]))
        for j in range(d):
            upd[j] = np.float32(upd[j] + np.float32(w * noise[t, j]))
    return upd


def test_fixed_point_sums_match_reference_host_loop(oracle_mod):
    rng = np.random.default_rng(5)
    m, d = 777, 5
    orc = oracle_mod.Oracle(k=1, m=m, n=8, d=d)
    scores = (rng.random(m) * 40 + 0.5).astype(np.float32)
    noise = (rng.standard_normal((m, d)) * 0.3).astype(np.float32)
    best = int(np.ceil(scores).min())
    sums = oracle_mod.update_sums(orc, best, scores, noise)
    assert sums[7] == 0 and sums[5] == 0 and sums[6] == 0
    u = np.array([oracle_mod.lib().orc_update_from_sum(int(sums[j]), 0) for j in range(d)], np.float32)
    ref = _host_loop_update(scores, noise, best)
    # FP tolerance of north_star (1e-5 relative), against the magnitude of the summed terms
    scale = np.abs(noise).sum(0) * best / np.ceil(scores).min()
    assert np.all(np.abs(u - ref) <= 1e-5 * scale)
    # order independence: any permutation of the samples gives identical integer sums
    perm = rng.permutation(m)
    assert np.array_equal(oracle_mod.update_sums(orc, best, scores[perm], noise[perm]), sums)
    # shards add up exactly
    a = oracle_mod.update_sums(orc, best, scores[:300], noise[:300])
    b = oracle_mod.update_sums(orc, best, scores[300:], noise[300:])
    assert np.array_equal(a + b, sums)


def test_fixed_point_poison(oracle_mod):
    orc = oracle_mod.Oracle(k=1, m=4, n=8, d=5)
    scores = np.array([0.0, 1.5, 2.0, 3.0], np.float32)  # ceilf(0) = 0 -> 0/0 = NaN, as the reference
    noise = np.ones((4, 5), np.float32)
    sums = oracle_mod.update_sums(orc, 0, scores, noise)
    assert sums[7] == 5
    assert np.isnan(oracle_mod.lib().orc_update_from_sum(0, int(sums[7])))


def test_oracle_weighted_plan_moves_trajectory(oracle_mod):
    k, m, n, d = 2, 64, 16, 5
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, sigma=0.05, update_mode=1, update_gain=1e-3,
                            collision_mode=oracle_mod.COLLISION_AS_SHIPPED)
    start, goal = synth.START, synth.GOALS[0]
    init = oracle_mod.plan_lerp(start, goal, n)
    r = orc.plan(start, init, 3)
    assert r["found_better"].all()
    # every waypoint of a climber moved by the same joint-space offset (noise is one d-vector
    # per sample, generate_trajectories.cu:139-140 subtracts u[j] from every waypoint)
    off = r["trajectories"] - init[None]
    assert np.abs(off).max() > 0
    assert np.allclose(off, off[:, :1], atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("k,m,n,d,ext,gain,coll", [
    (1, 64, 32, 5, True, 1e-3, 1),
    (3, 1000, 32, 5, False, 1e-3, 1),
    (2, 257, 50, 4, False, 1e-2, 0),
    (1, 4096, 64, 5, False, 10000.0, 1),   # the reference's literal gain (generate_trajectories.cu:140)
])
def test_weighted_plan_parity(capi, oracle_mod, k, m, n, d, ext, gain, coll):
    iters = 4
    kw = dict(sigma=0.05, update_mode=1, update_gain=gain, collision_mode=coll)
    ctx = capi.Context(k=k, m=m, n=n, d=d, **kw)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, **kw)
    g = synth.box_grid(64, 0.02)
    ctx.set_grid(g, 0.02)
    orc.set_grid(g, 0.02)
    start, goal = synth.START[:d], synth.GOALS[0][:d]
    init = capi.plan_lerp(start, goal, n)
    nz = synth.noise(iters * k, m, 1, d, 0.05).reshape(iters, k, m, 1, d) if ext else None
    a = ctx.plan(start, goal, init, iters, nz)
    b = orc.plan(start, init, iters, nz)
    st = ctx.plan_state()
    assert np.array_equal(st["trajectories"].view(np.uint32), b["trajectories"].view(np.uint32))
    assert np.array_equal(st["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert np.array_equal(st["found_better"], b["found_better"])
    assert a["best_climber"] == b["best_climber"]
    assert np.array_equal(a["best_traj"].view(np.uint32), b["best_traj"].view(np.uint32))
    a2 = ctx.plan(start, goal, init, iters, nz)   # graph replay is idempotent
    assert np.array_equal(a2["best_traj"], a["best_traj"])
    ctx.close()


@pytest.mark.gpu
def test_weighted_poison_gives_nan(capi, oracle_mod):
    # start == goal and zero noise sigma: every score is 0 -> 0/ceilf(0) = NaN, exactly what the
    # reference's formula does; both sides must poison the trajectory
    kw = dict(sigma=0.0, update_mode=1, update_gain=1.0, collision_mode=0)
    ctx = capi.Context(k=1, m=64, n=16, d=5, **kw)
    orc = oracle_mod.Oracle(k=1, m=64, n=16, d=5, **kw)
    start = np.zeros(5, np.float32)
    init = capi.plan_lerp(start, start, 16)
    a = ctx.plan(start, start, init, 2)
    b = orc.plan(start, init, 2)
    assert np.isnan(b["trajectories"]).all()
    assert np.isnan(ctx.plan_state()["trajectories"]).all()
    ctx.close()


@pytest.mark.gpu
def test_weighted_rejects_per_waypoint_noise(capi):
    with pytest.raises(capi.DoappError):
        capi.Context(k=1, m=64, n=16, d=5, n_noise=16, update_mode=1)
