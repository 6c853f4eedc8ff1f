"""CPU tests that PIN the oracle: against the reference's own code run on a B200
(tests/golden/ref_outputs.npz, produced by oracle/_ref/ref_harness), against the
reference's host loops restated in numpy (gt.test.cu:18-30,169-208), against
published known answers (Philox), and against stored KAT hashes."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "ref_outputs.npz"))


def f32(name):
    return GOLD[name].astype(np.uint32).view(np.float32)


# ---------------------------------------------------------------- reference goldens
def test_ref_matrix_golden():
    # perturbator/src/test_matrix.cu:7-24 — the one stored known answer in the reference
    assert list(GOLD["matmul_2x2_int"]) == [19, 22, 43, 50]
    a, b = f32("mat_a").reshape(-1, 3, 3), f32("mat_b").reshape(-1, 3, 3)
    ab = f32("mat_ab").reshape(-1, 3, 3)
    np.testing.assert_allclose(ab, np.einsum("tij,tjk->tik", a.astype(np.float64), b.astype(np.float64)),
                               rtol=0, atol=2e-6)
    # operator*(Matrix, Vector) as shipped sums each row and ignores v
    av = f32("mat_av_as_shipped").reshape(-1, 3)
    rows = np.zeros_like(av)
    for j in range(3):
        rows = (rows + a[:, :, j]).astype(np.float32)
    assert np.array_equal(av.view(np.uint32), rows.view(np.uint32))


def test_ref_kernel_one_iteration(oracle_mod):
    """optimize_trajectories (gt.cu:144-258) run for one iteration on a B200: the oracle in
    AS_SHIPPED mode must reproduce noisy trajectories, velocities and accelerations bit for bit
    from the reference's own noise, and smoothness/scores up to the float-atomicAdd order (D17)."""
    k, m, n, d = (int(v) for v in GOLD["kernel_shape_kmnd"])
    start, init = f32("kernel_start"), f32("kernel_initial_trajectory").reshape(n, d)
    noise = f32("kernel_noise").reshape(k, m, 1, d)
    noisy, vel, acc = (f32(x).reshape(k, m, n, d) for x in ("kernel_noisy", "kernel_vel", "kernel_acc"))
    smooth, scores = f32("kernel_smoothness").reshape(k, m), f32("kernel_scores").reshape(k, m + 1)
    assert 0.7 < noise.std() < 1.3  # curand_normal, unit variance (gt.cu:297-299)
    orc = oracle_mod.Oracle(k=1, m=1, n=n, d=d, collision_mode=oracle_mod.COLLISION_AS_SHIPPED,
                            sum_order=oracle_mod.SUM_SEQUENTIAL)
    for c in range(k):
        for t in range(m):
            r = orc.cost_sample(init, start, noise[c, t], details=True)
            assert np.array_equal(r["noisy"].view(np.uint32), noisy[c, t].view(np.uint32))
            assert np.array_equal(r["vel"].view(np.uint32), vel[c, t].view(np.uint32))
            assert np.array_equal(r["acc"].view(np.uint32), acc[c, t].view(np.uint32))
            assert r["collisions"] == 0
            np.testing.assert_allclose(r["smooth"], smooth[c, t], rtol=2e-6)
            np.testing.assert_allclose(r["score"], scores[c, t + 1], rtol=2e-6)
    # the same trajectory scored by 10 blocks differs in the last bits: the atomicAdd order (D17)
    assert np.ptp(scores[:, 0]) < 1e-5 * scores[0, 0]


@pytest.mark.parametrize("tag,n,res", [("100", 100, 0.01), ("64", 64, 0.02), ("128", 128, 0.01)])
def test_ref_grid_lookup(oracle_mod, tag, n, res):
    """CPUDistanceGrid ctor + operator() (cdg.cu:28-83,113-126) on a grid whose cells hold
    their logical index: oracle indices are identical wherever the reference is defined."""
    pts, vals = f32(f"grid_{tag}_points").reshape(-1, 3), f32(f"grid_{tag}_values")
    frame = f32(f"grid_{tag}_frame")
    assert frame[4] == 4 * n * n  # slice pitch = 4*L*W floats, logical index unaffected
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    orc.set_grid(np.zeros((n, n, n), np.uint8), res)
    idx = orc.grid_index(pts)
    ub = np.isnan(vals)  # coord == max: reference indexes out of bounds (D10), not evaluated
    expect = np.where(np.isfinite(vals), vals, -1).astype(np.int64)
    assert np.array_equal(idx[~ub], expect[~ub])
    assert (idx[ub] == -1).all()  # known deviation: half-open bounds
    assert (expect[~ub] >= 0).sum() > 1000 and ub.sum() > 10


# ------------------------------------------- the reference's host loops (gt.test.cu)
def _take_derivative(inp, start, dt):
    """gt.test.cu:18-30: backward difference, initial_waypoint as the prior of waypoint 0."""
    out = np.empty_like(inp)
    n, d = inp.shape
    for i in range(n):
        for j in range(d):
            prior = inp[i - 1, j] if i else start[j]
            out[i, j] = np.float32(np.float32(inp[i, j] - prior) / np.float32(dt))
    return out


def test_host_loops_of_reference_test(oracle_mod):
    """gt.test.cu:85-88 defaults (k=10, n=25, m=6, d=5, dt=1, start/goal in servo ticks) and
    its host recomputation :169-208 of noisy trajectories, derivatives, 0.5*sum(acc^2) and the
    update vector, against the oracle with smooth_scale 0.5 and the sequential sum order."""
    k, n, m, d, dt = 10, 25, 6, 5, 1.0
    start = np.array([1024, 1024, 1024, 0, 0], np.float32)
    rng = np.random.default_rng(467)
    traj = rng.standard_normal((k, n, d)).astype(np.float32)
    noise = rng.standard_normal((k, m, 1, d)).astype(np.float32)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, delta_t=dt, smooth_scale=0.5,
                            collision_mode=oracle_mod.COLLISION_AS_SHIPPED,
                            sum_order=oracle_mod.SUM_SEQUENTIAL)
    scores = np.empty((k, m), np.float32)
    for c in range(k):
        for t in range(m):
            noisy = (traj[c] + noise[c, t, 0]).astype(np.float32)          # :169-177
            vel = _take_derivative(noisy, start, dt)                        # :179-182
            acc = _take_derivative(vel, start, dt)                          # :183-185
            s = np.float32(0)
            for a in acc.ravel():                                           # :187-195
                s = np.float32(s + np.float32(a * a))
            scores[c, t] = np.float32(0.5) * s
            r = orc.cost_sample(traj[c], start, noise[c, t], details=True)
            assert np.array_equal(r["noisy"].view(np.uint32), noisy.view(np.uint32))
            assert np.array_equal(r["vel"].view(np.uint32), vel.view(np.uint32))
            assert np.array_equal(r["acc"].view(np.uint32), acc.view(np.uint32))
            assert np.float32(r["score"]) == scores[c, t]
    got = orc.cost_eval(traj, start, noise)
    assert np.array_equal(got["scores"].view(np.uint32), scores.view(np.uint32))
    # update vector, :200-208
    best = scores.ravel()[0]
    for s in scores.ravel()[1:]:
        if np.ceil(s) < np.ceil(best):
            best = s
    upd = np.zeros((k, d), np.float32)
    for c in range(k):
        for t in range(m):
            for j in range(d):
                w = np.float32(best / np.ceil(scores[c, t]))
                upd[c, j] = np.float32(upd[c, j] + np.float32(w * noise[c, t, 0, j]))
    assert np.array_equal(orc.update_vector(scores, noise).view(np.uint32), upd.view(np.uint32))


# ------------------------------------------------------------------ known answers
def test_philox_known_answers(oracle_mod):
    # Random123 v1.09 kat_vectors, philox4x32 10 rounds
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for ctr, key, out in kat:
        assert list(oracle_mod.philox(ctr, key)) == out


def test_noise_statistics_and_sharding(oracle_mod):
    orc = oracle_mod.Oracle(k=2, m=4096, n=8, sigma=0.5)
    z = orc.noise(3)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 0.5) < 0.01
    assert not np.array_equal(z, orc.noise(4))
    # a shard [1024, 2048) of m_global = 4096 draws exactly the global stream's slice
    sh = oracle_mod.Oracle(k=2, m=1024, n=8, sigma=0.5, sample_offset=1024, m_global=4096)
    assert np.array_equal(sh.noise(3), z[:, 1024:2048])


def test_sincos_accuracy(oracle_mod):
    x = np.concatenate([np.linspace(-7, 7, 50001), np.linspace(-1000, 1000, 20001),
                        [0.0, np.pi / 2, -np.pi / 2, np.pi, 1e5, -131071.9, 2e5, 1e7]]).astype(np.float32)
    s, c = oracle_mod.sincos(x)
    xd = x.astype(np.float64)
    small = np.abs(xd) < 1.4e5
    assert np.abs(s - np.sin(xd))[small].max() < 1.5e-7 and np.abs(c - np.cos(xd))[small].max() < 1.5e-7
    assert np.abs(s - np.sin(xd)).max() < 1e-3  # large arguments: bounded, deterministic
    assert np.all(np.isnan(oracle_mod.sincos(np.array([np.inf, np.nan], np.float32))[0]))
    lg = oracle_mod.logf(np.array([1.0, 0.5, 1e-3, 2.0 ** -24, 0.75], np.float32))
    np.testing.assert_allclose(lg, np.log([1.0, 0.5, 1e-3, 2.0 ** -24, 0.75]), rtol=3e-7, atol=1e-7)


# ------------------------------------------------------------------ FK and the grid
def test_fk_zero_pose_and_d4(oracle_mod):
    p = oracle_mod.fk_points(np.zeros((1, 5), np.float32))[0]
    # link offsets of gt.cu:407,418,429,440 (= rexarm.urdf joint origins) accumulate along z
    np.testing.assert_allclose(p[:4, 2], [0.0715, 0.1165, 0.2165, 0.3165], atol=1e-6)
    np.testing.assert_allclose(p[:4, 1], -0.0015, atol=1e-7)
    # after wrist_rotate: Rrpy(0,-pi/2,0) maps z -> -x, then t = (0,-0.005,0.07)   gt.cu:456-459
    np.testing.assert_allclose(p[4], [-0.3965, -0.019, 0.07], atol=1e-6)
    rng = np.random.default_rng(3)
    wp = rng.uniform(-3, 3, (200, 5)).astype(np.float32)
    wp[:, 4] = 0
    a, b = oracle_mod.fk_points(wp), oracle_mod.fk_points(wp[:, :4].copy())
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))  # 4-joint waypoint: theta4 := 0 (D3)
    # with the matrix.cuh mat-vec as shipped every point is a row sum of R(theta): offsets vanish
    s = oracle_mod.fk_points(wp, as_shipped_matvec=True)
    assert np.abs(s).max() <= 3.0 and not np.allclose(s, a)


def test_occupancy_frame_matches_og_cpp(oracle_mod):
    """og.cpp:26,29-40: index = size_t((c + size/2)/granularity) in double, flat ix + N*iy + N^2*iz."""
    n, gran = 50, 0.04
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    orc.set_grid(np.zeros((n, n, n), np.uint8), gran, oracle_mod.FRAME_OG)
    rng = np.random.default_rng(4)
    pts = rng.uniform(-1.2, 1.2, (4000, 3)).astype(np.float32)
    half = n * gran / 2
    exp = []
    for x, y, z in pts.astype(np.float64):
        q = [(x + half) / gran, (y + half) / gran, (z + half) / gran]
        if min(q) < 0 or max(q) >= n:
            exp.append(-1)
        else:
            ix, iy, iz = (int(v) for v in q)
            exp.append(ix + n * iy + n * n * iz)
    assert np.array_equal(orc.grid_index(pts), np.array(exp, np.int64))


def test_scatter_is_set_per_point(oracle_mod):
    n, res = 64, 0.02
    orc = oracle_mod.Oracle(k=1, m=1, n=1)
    orc.set_grid(np.zeros((n, n, n), np.uint8), res)
    pts = synth.cloud(20000, 0)
    inside = orc.scatter(pts, 1, True)
    idx = orc.grid_index(pts)
    exp = np.zeros(n ** 3, np.uint8)
    exp[idx[idx >= 0]] = 1
    assert inside == (idx >= 0).sum() and np.array_equal(orc.cells.ravel(), exp)
    orc.scatter(pts[:10], 5, False)  # no clear: values accumulate
    assert orc.cells.ravel()[idx[:10][idx[:10] >= 0]].tolist() == [5] * int((idx[:10] >= 0).sum())
    assert orc.scatter(np.zeros((0, 3), np.float32), 1, True) == 0 and orc.cells.sum() == 0


# ------------------------------------------------------------------ planner loop
def test_plan_keep_best_properties(oracle_mod):
    k, m, n = 3, 64, 32
    orc = oracle_mod.Oracle(k=k, m=m, n=n, sigma=0.05, select_mode=oracle_mod.SELECT_FLOAT)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    start, goal = synth.START, synth.GOALS[1]
    init = oracle_mod.plan_lerp(start, goal, n)
    np.testing.assert_allclose(init[-1], goal, atol=1e-6)
    s0, _ = orc.cost_sample(init, start, np.zeros((1, 5), np.float32))
    prev = s0
    for iters in (0, 1, 4, 8):
        r = orc.plan(start, init, iters)
        assert r["best_score"] <= prev  # keep-best never gets worse, more iterations never hurt
        prev = r["best_score"]
        assert (r["best_climber"] == -1) == (not r["found_better"].any() or
                                             not r["found_better"][int(np.argmin(r["scores"]))])
    assert prev < s0
    # threads do not change anything; sharded evaluation covers the same samples
    a, b = orc.plan(start, init, 4, threads=1), orc.plan(start, init, 4, threads=5)
    assert np.array_equal(a["trajectories"], b["trajectories"])
    # velocities are backward differences of the published trajectory (ps.cu:129-166)
    v = np.empty_like(a["best_traj"])
    v[0] = (a["best_traj"][0] - start) / np.float32(0.1)
    v[1:] = (a["best_traj"][1:] - a["best_traj"][:-1]) / np.float32(0.1)
    assert np.array_equal(v.view(np.uint32), a["best_vel"].view(np.uint32))


def test_trajectory_message_layout(oracle_mod):
    """ps.cu:179-209: n+2 points — current state, n waypoints stamped dt*i, goal with zero velocity."""
    n, d = 6, 5
    rng = np.random.default_rng(9)
    cur, goal = rng.standard_normal(d).astype(np.float32), rng.standard_normal(d).astype(np.float32)
    traj, vel = rng.standard_normal((n, d)).astype(np.float32), rng.standard_normal((n, d)).astype(np.float32)
    pos, velo, tfs = oracle_mod.build_message(cur, goal, traj, vel, 0.1)
    assert pos.shape == (n + 2, d)
    assert np.array_equal(pos[0], cur) and np.array_equal(pos[1:-1], traj) and np.array_equal(pos[-1], goal)
    np.testing.assert_allclose(velo[0], (traj[0] - cur) / 0.1, rtol=1e-6)
    assert np.array_equal(velo[1:n], vel[1:]) and not velo[n].any() and not velo[-1].any()
    np.testing.assert_allclose(tfs[:-1], [0] + [0.1 * i for i in range(n)], atol=1e-7)


# ------------------------------------------------------------------ stored KAT hashes
def _c1(oracle_mod):
    """BASELINE config 1: 64^3 grid (2 cm), one box obstacle, 32 waypoints x 64 samples."""
    orc = oracle_mod.Oracle(k=1, m=64, n=32, sigma=0.1)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    base = oracle_mod.plan_lerp(synth.START, synth.GOALS[0], 32)[None]
    return orc, base


def test_c1_known_answers(oracle_mod):
    orc, base = _c1(oracle_mod)
    r = orc.cost_eval(base, synth.START, None, iteration=0)
    got = {"scores_sha256": hashlib.sha256(r["scores"].tobytes()).hexdigest(),
           "collisions_sha256": hashlib.sha256(r["collisions"].tobytes()).hexdigest(),
           "noise_sha256": hashlib.sha256(orc.noise(0).tobytes()).hexdigest(),
           "argmin": int(r["argmin"]), "min_score_bits": int(np.float32(r["min_score"]).view(np.uint32)),
           "collisions_total": int(r["collisions"].sum())}
    path = os.path.join(HERE, "golden", "c1_kat.json")
    if os.environ.get("DOAPP_WRITE_GOLDEN"):
        json.dump(got, open(path, "w"), indent=1)
    assert got == json.load(open(path))
