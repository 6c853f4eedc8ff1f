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
