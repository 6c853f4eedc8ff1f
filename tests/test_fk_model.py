"""URDF-correct forward kinematics with per-link collision spheres:
rexarm_description/robots/rexarm.urdf:61-66,80-85,106-111,125-130,144-149 joint origins and
axes, accumulated link frames, spheres along the link boxes.  CPU tests pin the oracle's chain
against an independent fp64 homogeneous-transform product; GPU tests pin the CUDA path to the
oracle bit for bit (sphere centres, collision counts, scores, planner results)."""
import ctypes as C

import numpy as np
import pytest

from tests import synth

AXES = {0: np.array([1.0, 0, 0]), 1: np.array([0, 1.0, 0]), 2: np.array([0, 0, 1.0])}


def _rot(axis, th):
    """Rodrigues, fp64."""
    a = AXES[axis]
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def _fk64(model, wp):
    """T_i = T_{i-1} Trans(origin_i) Rot(axis_i, theta_i) with 4x4 matrices in fp64."""
    out = np.zeros((model.num_spheres, 3))
    T = np.eye(4)
    for link in range(5):
        step = np.eye(4)
        step[:3, 3] = np.array(list(model.joint_origin[link]), np.float64)
        step[:3, :3] = _rot(model.joint_axis[link], float(wp[link]) if link < len(wp) else 0.0)
        T = T @ step
        for q in range(model.num_spheres):
            sp = model.spheres[q]
            if sp.link == link:
                out[q] = (T @ np.array([*sp.centre, 1.0]))[:3]
    return out


def _custom_model(mod):
    """x-axis joint, shuffled link order, all 16 spheres, a zero radius."""
    m = mod.fk_model_rexarm()
    rng = np.random.default_rng(3)
    m.joint_axis[1] = 0
    m.joint_axis[3] = 0
    m.joint_origin[2][0] = 0.02
    m.num_spheres = 16
    links = rng.permutation(np.arange(16) % 5)
    for q in range(16):
        m.spheres[q].link = int(links[q])
        for a in range(3):
            m.spheres[q].centre[a] = float(np.float32(rng.uniform(-0.05, 0.1)))
        m.spheres[q].radius = float(np.float32(rng.uniform(0.0, 0.03))) if q else 0.0
    return m


def _copy_model(src, dst_cls):
    """oracle FkModel <-> capi FkModel (the ABI one carries struct_size)."""
    dst = dst_cls()
    if hasattr(dst, "struct_size"):
        dst.struct_size = C.sizeof(dst_cls)
    for link in range(5):
        dst.joint_axis[link] = src.joint_axis[link]
        for a in range(3):
            dst.joint_origin[link][a] = src.joint_origin[link][a]
    dst.num_spheres = src.num_spheres
    for q in range(src.num_spheres):
        dst.spheres[q].link = src.spheres[q].link
        dst.spheres[q].radius = src.spheres[q].radius
        for a in range(3):
            dst.spheres[q].centre[a] = src.spheres[q].centre[a]
    return dst


# ------------------------------------------------------------------ CPU: the oracle

def test_oracle_zero_pose_matches_urdf_origins(oracle_mod):
    m = oracle_mod.fk_model_rexarm()
    assert m.num_spheres == 13 and list(m.joint_axis) == [2, 1, 1, 1, 2]
    pts = oracle_mod.fk_model_points(m, np.zeros((1, 5), np.float32))[0]
    # gripper_center: the joint origins of rexarm.urdf:64,83,109,128,147 plus :197, summed
    np.testing.assert_allclose(pts[12], [0.0, -0.0015 - 0.0125 - 0.005,
                                         0.0715 + 0.045 + 0.1 + 0.1 + 0.08 + 0.07], atol=1e-7)
    # base yaw turns everything about z; the gripper keeps its height
    yaw = oracle_mod.fk_model_points(m, np.array([[np.pi / 2, 0, 0, 0, 0]], np.float32))[0]
    np.testing.assert_allclose(yaw[:, 2], pts[:, 2], atol=1e-7)
    # shoulder pitch of +90 deg about y lays the arm along +x at the shoulder's height
    lay = oracle_mod.fk_model_points(m, np.array([[0, np.pi / 2, 0, 0, 0]], np.float32))[0]
    np.testing.assert_allclose(lay[12], [0.1 + 0.1 + 0.08 + 0.07, -0.019, 0.0715 + 0.045], atol=1e-6)


@pytest.mark.parametrize("d", [5, 4])
def test_oracle_chain_matches_fp64_transforms(oracle_mod, d):
    rng = np.random.default_rng(11)
    wp = rng.uniform(-3.2, 3.2, (300, d)).astype(np.float32)
    for m in (oracle_mod.fk_model_rexarm(), _custom_model(oracle_mod)):
        got = oracle_mod.fk_model_points(m, wp)
        ref = np.stack([_fk64(m, w) for w in wp])
        np.testing.assert_allclose(got, ref, atol=2e-6, rtol=0)


def test_oracle_chain_differs_from_reference_chain(oracle_mod):
    """D14: the reference's point chain is not the URDF's.  Up to the wrist they agree at the
    zero pose (both just add the joint origins); the reference then turns the POINT by the
    gripper_center joint's fixed pitch (gt.cu:456-459) instead of the frame."""
    m = oracle_mod.fk_model_rexarm()
    z = np.zeros((1, 5), np.float32)
    ref0 = oracle_mod.fk_points(z)[0]
    np.testing.assert_allclose(ref0[3], [0, -0.0015, 0.0715 + 0.045 + 0.1 + 0.1], atol=1e-6)
    np.testing.assert_allclose(ref0[4], [-(0.3165 + 0.08), -0.019, 0.07], atol=1e-6)
    assert not np.allclose(oracle_mod.fk_model_points(m, z)[0][12], ref0[4], atol=1e-3)
    w = np.array([[0.7, 0.3, -0.4, 0.5, 0.2]], np.float32)
    assert not np.allclose(oracle_mod.fk_model_points(m, w)[0][12], oracle_mod.fk_points(w)[0][4], atol=1e-3)


def test_oracle_sphere_collisions(oracle_mod):
    """U8: a sphere collides iff the cell at its centre is set; F32: iff the distance cell is
    <= radius + collision_distance."""
    m = oracle_mod.fk_model_rexarm()
    orc = oracle_mod.Oracle(k=1, m=1, n=1, fk_model=m)
    wp = np.zeros((1, 5), np.float32)
    cen = oracle_mod.fk_model_points(m, wp)[0]
    n, res = 64, 0.02
    cells = np.zeros((n, n, n), np.uint8)
    orc.set_grid(cells, res)
    hit = orc.grid_index(cen[[0, 5, 12]])
    assert (hit >= 0).all()
    cells.ravel()[hit] = 1
    orc.set_grid(cells, res)
    _, coll = orc.cost_sample(wp, wp[0], np.zeros((1, 5), np.float32))
    marked = set(hit.tolist())
    assert coll == sum(1 for q in range(13) if int(orc.grid_index(cen[q:q + 1])[0]) in marked)
    assert 3 <= coll < 13
    dist = np.full((n, n, n), 0.06, np.float32)  # > collision_distance, <= 0.05 + radius for r >= 0.01
    orc.set_grid(dist, res)
    _, coll = orc.cost_sample(wp, wp[0], np.zeros((1, 5), np.float32))
    assert coll == 13
    dist[:] = 0.0651  # only spheres with radius >= 0.0151 still reach
    orc.set_grid(dist, res)
    _, coll = orc.cost_sample(wp, wp[0], np.zeros((1, 5), np.float32))
    assert coll == sum(1 for q in range(13) if np.float32(m.spheres[q].radius) + np.float32(0.05) >= np.float32(0.0651))
    assert 0 < coll < 13


def test_abi_default_model_equals_oracle_model(capi, oracle_mod):
    a, b = capi.fk_model_rexarm(), oracle_mod.fk_model_rexarm()
    assert a.struct_size == C.sizeof(capi.FkModel) and a.num_spheres == b.num_spheres
    assert list(a.joint_axis) == list(b.joint_axis)
    for link in range(5):
        assert np.array_equal(np.array(list(a.joint_origin[link]), np.float32).view(np.uint32),
                              np.array(list(b.joint_origin[link]), np.float32).view(np.uint32))
    for q in range(a.num_spheres):
        assert a.spheres[q].link == b.spheres[q].link
        got = np.array([*a.spheres[q].centre, a.spheres[q].radius], np.float32)
        ref = np.array([*b.spheres[q].centre, b.spheres[q].radius], np.float32)
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), q


# ------------------------------------------------------------------ GPU: parity

@pytest.mark.gpu
@pytest.mark.parametrize("d", [5, 4])
@pytest.mark.parametrize("custom", [False, True])
def test_model_points_bit_exact(capi, oracle_mod, d, custom):
    rng = np.random.default_rng(21)
    wp = np.concatenate([rng.uniform(-3.2, 3.2, (4096, d)), rng.uniform(-2e5, 2e5, (64, d)),
                         np.zeros((1, d)), -np.zeros((1, d))]).astype(np.float32)
    om = _custom_model(oracle_mod) if custom else oracle_mod.fk_model_rexarm()
    with capi.Context(k=1, m=1, n=1, d=d) as ctx:
        with pytest.raises(capi.DoappError):
            ctx._fk_spheres = 13
            ctx.debug_fk_model(wp)  # no model set yet
        ctx.set_fk_model(_copy_model(om, capi.FkModel))
        got = ctx.debug_fk_model(wp)
    ref = oracle_mod.fk_model_points(om, wp)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


MODEL_CASES = [  # (k, m, n, d, n_noise, grid, frame, sigma, custom)
    (1, 64, 32, 5, 1, "u8", 0, 0.1, False),
    (1, 64, 32, 5, 1, "f32", 0, 0.1, False),
    (3, 37, 50, 4, 1, "u8", 0, 0.3, True),
    (2, 33, 33, 5, 33, "f32", 0, 0.2, True),     # per-waypoint noise
    (1, 1024, 64, 5, 1, "u8", 0, 0.5, False),
    (2, 40, 48, 5, 1, "u8", 1, 0.3, False),      # OG frame -> generic grid kernel
    (1, 9, 700, 5, 1, "u8", 0, 0.05, False),     # many threads per sample, cluster geometry
    (1, 64, 32, 5, 1, "u8", 0, 3.0e5, False),    # angles beyond the fast sin/cos range
]


@pytest.mark.gpu
@pytest.mark.parametrize("k,m,n,d,n_noise,grid,frame,sigma,custom", MODEL_CASES)
def test_model_cost_parity(capi, oracle_mod, k, m, n, d, n_noise, grid, frame, sigma, custom):
    om = _custom_model(oracle_mod) if custom else oracle_mod.fk_model_rexarm()
    ctx = capi.Context(k=k, m=m, n=n, d=d, n_noise=n_noise, sigma=sigma)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, n_noise=n_noise, sigma=sigma, fk_model=om)
    g = synth.box_grid(64, 0.02) if grid == "u8" else synth.box_distance_grid(64, 0.02)
    if frame == 1:  # the box sits outside the arm's reach in the z-centred OG frame
        g = (np.random.default_rng(8).random((64, 64, 64)) < 0.1).astype(np.uint8)
    ctx.set_grid(g, 0.02, frame)
    orc.set_grid(g, 0.02, frame)
    start = synth.START[:d]
    base = np.stack([capi.plan_lerp(start, synth.GOALS[c % 2][:d], n) for c in range(k)])
    nz = synth.noise(k, m, n_noise, d, sigma)
    ref_chain = ctx.cost_eval(base, start, nz, iteration=3)
    ctx.set_fk_model(_copy_model(om, capi.FkModel))
    for noise in (nz, None):
        a = ctx.cost_eval(base, start, noise, iteration=3)
        b = orc.cost_eval(base, start, noise, iteration=3)
        assert np.array_equal(a["collisions"], b["collisions"])
        assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
        assert a["argmin"] == b["argmin"] and a["min_score"] == b["min_score"]
    if sigma < 10:
        assert b["collisions"].sum() > 0
    # back to the reference's chain: same bits as before the model was set
    ctx.set_fk_model(None)
    again = ctx.cost_eval(base, start, nz, iteration=3)
    assert np.array_equal(again["scores"].view(np.uint32), ref_chain["scores"].view(np.uint32))
    assert np.array_equal(again["collisions"], ref_chain["collisions"])
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("update_mode,persist", [(0, "1"), (0, "0"), (1, "1")])
def test_model_plan_parity(capi, oracle_mod, monkeypatch, update_mode, persist):
    """Planner loop on the FK model: the persistent one-launch planner (keep-best), the
    launch-per-iteration graph, and the weighted update rule."""
    monkeypatch.setenv("DOAPP_PERSIST", persist)
    k, m, n, iters = 2, 96, 48, 6
    om = oracle_mod.fk_model_rexarm()
    kw = dict(sigma=0.05, select_mode=1, update_mode=update_mode, update_gain=1e-3)
    ctx = capi.Context(k=k, m=m, n=n, **kw)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, fk_model=om, **kw)
    g = synth.box_grid(64, 0.02)
    ctx.set_grid(g, 0.02)
    orc.set_grid(g, 0.02)
    init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
    plain = ctx.plan(synth.START, synth.GOALS[1], init, iters)["best_traj"].copy()
    ctx.set_fk_model(capi.fk_model_rexarm())
    a = ctx.plan(synth.START, synth.GOALS[1], init, iters)
    assert (ctx.plan_mode() & 1) == (1 if update_mode == 0 and persist == "1" else 0)
    b = orc.plan(synth.START, init, iters)
    st = ctx.plan_state()
    assert np.array_equal(st["trajectories"].view(np.uint32), b["trajectories"].view(np.uint32))
    assert np.array_equal(st["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert np.array_equal(st["found_better"], b["found_better"])
    assert np.array_equal(a["best_traj"].view(np.uint32), b["best_traj"].view(np.uint32))
    assert np.float32(a["best_score"]) == np.float32(b["best_score"])
    # replay through the cached graph, then without the model
    a2 = ctx.plan(synth.START, synth.GOALS[1], init, iters)
    assert np.array_equal(a2["best_traj"], a["best_traj"])
    ctx.set_fk_model(None)
    assert np.array_equal(ctx.plan(synth.START, synth.GOALS[1], init, iters)["best_traj"], plain)
    ctx.close()


@pytest.mark.gpu
def test_model_rejects_bad_input(capi):
    with capi.Context(k=1, m=1, n=1) as ctx:
        m = capi.fk_model_rexarm()
        m.struct_size = 8
        with pytest.raises(capi.DoappError):
            ctx.set_fk_model(m)
        for field, val in (("num_spheres", 0), ("num_spheres", 17)):
            m = capi.fk_model_rexarm()
            setattr(m, field, val)
            with pytest.raises(capi.DoappError):
                ctx.set_fk_model(m)
        m = capi.fk_model_rexarm()
        m.joint_axis[2] = 3
        with pytest.raises(capi.DoappError):
            ctx.set_fk_model(m)
        m = capi.fk_model_rexarm()
        m.spheres[4].link = 5
        with pytest.raises(capi.DoappError):
            ctx.set_fk_model(m)
