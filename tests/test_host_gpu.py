"""The C++ PlanningState mirror (doapp_b200/host/) driven like perturbator/src/main.cu drives the
reference: joint_state + goal_state in, plan(), publish_trajectory() -> JointTrajectory out.
The published message must equal the oracle's plan + message assembly (ps.cu:179-209)."""
import json
import os
import subprocess

import numpy as np
import pytest

from tests import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "doapp_b200", "lib", "planning_state_test")


def test_host_library_built(capi):
    assert os.path.exists(BIN)
    lib = os.path.join(ROOT, "doapp_b200", "lib", "libdoapp_planning_state.so")
    syms = subprocess.run(["nm", "-DC", lib], capture_output=True, text=True, check=True).stdout
    for name in ("PlanningState::PlanningState", "PlanningState::plan()", "PlanningState::plan_lerp",
                 "PlanningState::publish_trajectory", "PlanningState::updateArmState",
                 "PlanningState::updateGoalState", "PlanningState::updatePointCloud",
                 "PlanningState::updateDepthFrames", "PlanningState::buildDistanceField"):
        assert name in syms, name


@pytest.mark.gpu
@pytest.mark.parametrize("k,m,n,d,iters", [(4, 64, 32, 5, 6), (2, 33, 50, 4, 3)])
def test_planning_state_publishes_oracle_trajectory(capi, oracle_mod, tmp_path, k, m, n, d, iters):
    gn, res = 64, 0.02
    cloud = synth.cloud(30000, 0)
    path = tmp_path / "cloud.bin"
    cloud.astype(np.float32).tofile(path)
    out = subprocess.run([BIN, str(k), str(m), str(n), str(d), str(gn), str(res), str(iters), str(path)],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    msg = json.loads(out.stdout)
    pos = np.array([p["positions"] for p in msg["points"]], np.float32)
    vel = np.array([p["velocities"] for p in msg["points"]], np.float32)
    tfs = np.array([p["t"] for p in msg["points"]], np.float32)

    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, select_mode=oracle_mod.SELECT_CEIL_INT)
    orc.set_grid(np.zeros((gn, gn, gn), np.uint8), res)
    orc.scatter(cloud, 1, True)
    cur = np.zeros(d, np.float32)
    goal = np.array([0.0, -1.0, 1.5, -1.5, 0.0], np.float32)[:d]
    init = oracle_mod.plan_lerp(cur, goal, n)
    r = orc.plan(cur, init, iters)
    epos, evel, etfs = oracle_mod.build_message(cur, goal, r["best_traj"], r["best_vel"], 0.1)
    assert pos.shape == (n + 2, d)
    assert np.array_equal(pos.view(np.uint32), epos.view(np.uint32))
    assert np.array_equal(vel.view(np.uint32), evel.view(np.uint32))
    np.testing.assert_allclose(tfs, etfs, rtol=1e-6, atol=1e-7)
    assert np.float32(msg["cost"]) == np.float32(r["best_score"])


@pytest.mark.gpu
def test_follower_padding_and_distance_field(capi, oracle_mod, tmp_path):
    """Row 4 of SURVEY 8f: points padded to the 6 entries the follower reads
    (trajectory_follower.cpp:176-178), and the plan on the derived distance grid (gt.cu:412)."""
    k, m, n, d, iters, gn, res = 2, 64, 32, 5, 4, 64, 0.02
    cloud = synth.cloud(30000, 0)
    path = tmp_path / "cloud.bin"
    cloud.astype(np.float32).tofile(path)
    out = subprocess.run([BIN, str(k), str(m), str(n), str(d), str(gn), str(res), str(iters), str(path),
                          "6", "edt"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    msg = json.loads(out.stdout)
    pos = np.array([p["positions"] for p in msg["points"]], np.float32)
    vel = np.array([p["velocities"] for p in msg["points"]], np.float32)
    assert pos.shape == (n + 2, 6) and vel.shape == (n + 2, 6)
    assert np.all(pos[:, 5] == np.float32(0.7)) and np.all(vel[:, 5] == 0)  # sixth servo held
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, select_mode=oracle_mod.SELECT_CEIL_INT)
    orc.set_grid(np.zeros((gn, gn, gn), np.uint8), res)
    orc.scatter(cloud, 1, True)
    orc.set_grid(orc.distance_field(), res)
    cur = np.zeros(d, np.float32)
    goal = np.array([0.0, -1.0, 1.5, -1.5, 0.0], np.float32)
    r = orc.plan(cur, oracle_mod.plan_lerp(cur, goal, n), iters)
    epos, evel, _ = oracle_mod.build_message(cur, goal, r["best_traj"], r["best_vel"], 0.1)
    assert np.array_equal(pos[:, :d].view(np.uint32), epos.view(np.uint32))
    assert np.array_equal(vel[:, :d].view(np.uint32), evel.view(np.uint32))
    assert np.float32(msg["cost"]) == np.float32(r["best_score"])
    # the consumer's side (trajectory_follower.cpp:148-215 replayed at 100 Hz): the message is
    # well formed for 6 servos, ends on the goal, and its velocities are the finite differences
    times = [p["t"] for p in msg["points"]]
    for key, intended in (("follower_as_shipped", False), ("follower_intended", True)):
        f = msg[key]
        ticks, now = _replay_ticks(times, intended)
        assert f["well_formed"] == 1 and f["ticks"] == ticks and abs(f["duration"] - now) < 1e-6
        assert f["max_interp_gap"] <= 1e-12 and f["max_velocity_mismatch"] < 1e-4
        assert np.array_equal(np.array(f["last_command"], np.float32), pos[-1])
        step = np.abs(np.diff(pos.astype(np.float64), axis=0)).max()
        assert f["max_command_step"] <= step + 1e-12
    # as shipped the segment end is the previous point's stamp: the commanded point runs up to
    # one segment (delta_t) ahead of the clock; with the intended end it never leads
    assert 0.08 < msg["follower_as_shipped"]["max_lead"] < 0.1 + 1e-9
    assert msg["follower_intended"]["max_lead"] <= 1e-12
    assert 0 <= msg["follower_intended"]["ticks"] - msg["follower_as_shipped"]["ticks"] <= 2


def _replay_ticks(t, intended, tick=0.01):
    """trajectory_follower.cpp:139-144,157-160,206-211: index stepping only."""
    prev, nxt, now, ticks = 0, (1 if len(t) > 1 else 0), 0.0, 0
    while True:
        ticks += 1
        if prev + 1 == len(t):
            break
        if now > (t[nxt] if intended else t[prev]):
            prev = nxt
            if nxt + 1 < len(t):
                nxt += 1
        now += tick
    return ticks, now


@pytest.mark.gpu
def test_urdf_kinematics_through_planning_state(capi, oracle_mod, tmp_path):
    """PlanningState::useUrdfKinematics: the whole host path on the URDF chain + spheres
     equals the oracle with the same FK model."""
    k, m, n, d, iters, gn, res = 1, 128, 32, 5, 5, 64, 0.02
    cloud = synth.cloud(30000, 0)
    path = tmp_path / "cloud.bin"
    cloud.astype(np.float32).tofile(path)
    out = subprocess.run([BIN, str(k), str(m), str(n), str(d), str(gn), str(res), str(iters), str(path),
                          "5", "-", "urdf"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    msg = json.loads(out.stdout)
    pos = np.array([p["positions"] for p in msg["points"]], np.float32)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, d=d, select_mode=oracle_mod.SELECT_CEIL_INT,
                            fk_model=oracle_mod.fk_model_rexarm())
    orc.set_grid(np.zeros((gn, gn, gn), np.uint8), res)
    orc.scatter(cloud, 1, True)
    cur = np.zeros(d, np.float32)
    goal = np.array([0.0, -1.0, 1.5, -1.5, 0.0], np.float32)
    r = orc.plan(cur, oracle_mod.plan_lerp(cur, goal, n), iters)
    epos, _, _ = oracle_mod.build_message(cur, goal, r["best_traj"], r["best_vel"], 0.1)
    assert np.array_equal(pos.view(np.uint32), epos.view(np.uint32))
    assert np.float32(msg["cost"]) == np.float32(r["best_score"])


@pytest.mark.gpu
def test_update_distance_grid_from_kdtree(capi, tmp_path):
    """PlanningState::updateDistanceGrid(const KDTree&) (planning_state.cuh:25-28, nanoflann tree
    over the obstacle points) gives the plan of the explicit scatter + distance-field calls."""
    k, m, n, d, iters, gn, res = 2, 64, 32, 5, 4, 64, 0.02
    path = tmp_path / "cloud.bin"
    synth.cloud(30000, 0).astype(np.float32).tofile(path)
    outs = []
    for mode in ("edt", "kdtree"):
        out = subprocess.run([BIN, str(k), str(m), str(n), str(d), str(gn), str(res), str(iters),
                              str(path), "6", mode], capture_output=True, text=True, timeout=120)
        if mode == "kdtree" and out.returncode == 7:
            pytest.skip("host library built without nanoflann")
        assert out.returncode == 0, out.stderr
        outs.append(json.loads(out.stdout))
    assert outs[0]["points"] == outs[1]["points"] and outs[0]["cost"] == outs[1]["cost"]


@pytest.mark.gpu
def test_motion_planner_node_loop(capi):
    """The node's plan / publish cycle of perturbator/src/main.cu:31-44 at its shipped shape
    (k = 1000, n = 50, d = 4, m = 5, 100^3 grid at 1 cm), driven without ROS."""
    exe = os.path.join(os.path.dirname(BIN), "motion_planner_main")
    out = subprocess.run([exe, "3", "10"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("cycle")]
    assert len(lines) == 3 and all("52 points" in ln for ln in lines)  # n + 2 points per message
