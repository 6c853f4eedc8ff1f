"""Long-run stress of the hand-rolled synchronisation (VERDICT r01 "weak #4"): 100 000 optimiser
iterations in ONE persistent launch (tagged key words, deep speculation, version restarts) must
leave exactly the state of 100 000 separate cost launches, for several seeds and both
generations of the persistent planner."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def _plan(capi, monkeypatch, env, seed, iters, k=4, m=256, n=32):
    for key in ("DOAPP_PERSIST", "DOAPP_PERSIST_V", "DOAPP_NO_GRAPH", "DOAPP_PERSIST_DEPTH"):
        monkeypatch.delenv(key, raising=False)
    for key, val in env.items():
        monkeypatch.setenv(key, val)
    with capi.Context(k=k, m=m, n=n, sigma=0.02, select_mode=1, seed=seed) as ctx:
        ctx.set_grid(synth.box_grid(64, 0.02), 0.02)
        init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
        r = ctx.plan(synth.START, synth.GOALS[1], init, iters)
        st = ctx.plan_state()
        return r, st, ctx.plan_mode(), ctx.plan_stats()


@pytest.mark.parametrize("seed", [1234, 7, 20191])
def test_100k_iterations_persistent_equals_launch_per_iteration(capi, monkeypatch, seed):
    iters = 100_000
    ref = _plan(capi, monkeypatch, {"DOAPP_PERSIST": "0", "DOAPP_NO_GRAPH": "1"}, seed, iters)
    assert ref[2] & 1 == 0
    for env in ({}, {"DOAPP_PERSIST_DEPTH": "15"}, {"DOAPP_PERSIST_V": "1"}):
        got = _plan(capi, monkeypatch, env, seed, iters)
        assert got[2] & 1 == 1, "the persistent planner must run"
        assert np.array_equal(ref[1]["trajectories"].view(np.uint32), got[1]["trajectories"].view(np.uint32))
        assert np.array_equal(ref[1]["scores"].view(np.uint32), got[1]["scores"].view(np.uint32))
        assert np.array_equal(ref[0]["best_traj"].view(np.uint32), got[0]["best_traj"].view(np.uint32))
        assert np.array_equal(ref[1]["found_better"], got[1]["found_better"])
    assert ref[1]["found_better"].any()
