"""plan_pipeline_kernel (second-generation persistent planner) on the paths the other planner
tests do not reach: the joint-limit term, and the rare IEEE fallback taken when an angle lies
outside the ranges the straight-line arithmetic was verified for (denormal, exactly zero,
>= 2^17) -- with externally supplied noise so the odd angles are planted exactly.  Every case is
compared with the oracle and with the launch-per-iteration path (DOAPP_PERSIST=0)."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def _run(capi, monkeypatch, persist, k, m, n, iters, noise, start, init, limits, **kw):
    for key in ("DOAPP_PERSIST", "DOAPP_PERSIST_V", "DOAPP_PERSIST_DEPTH"):
        monkeypatch.delenv(key, raising=False)
    for key, val in persist.items():
        monkeypatch.setenv(key, val)
    with capi.Context(k=k, m=m, n=n, **kw) as ctx:
        if limits is not None:
            ctx.set_limits(*limits)
        ctx.set_grid(synth.box_grid(64, 0.02), 0.02)
        r = ctx.plan(start, synth.GOALS[1], init, iters, noise)
        st = ctx.plan_state()
        return r, st, ctx.plan_mode()


def _same(a, b):
    assert np.array_equal(a["trajectories"].view(np.uint32), b["trajectories"].view(np.uint32))
    assert np.array_equal(a["scores"].view(np.uint32), b["scores"].view(np.uint32))
    assert np.array_equal(a["found_better"], b["found_better"])


@pytest.mark.parametrize("k,m,n", [(1, 300, 64), (3, 96, 33)])
def test_pipeline_planner_with_joint_limits(capi, oracle_mod, monkeypatch, k, m, n):
    iters = 12
    lo, hi = -0.3 * np.ones(5, np.float32), 0.6 * np.ones(5, np.float32)
    kw = dict(sigma=0.2, select_mode=1, limit_weight=2.0)
    start = synth.START
    init = capi.plan_lerp(start, synth.GOALS[1], n)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, **kw)
    orc.set_limits(lo, hi)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    want = orc.plan(start, init, iters, threads=8)
    for persist in ({}, {"DOAPP_PERSIST_DEPTH": "2"}, {"DOAPP_PERSIST": "0"}):
        r, st, mode = _run(capi, monkeypatch, persist, k, m, n, iters, None, start, init, (lo, hi), **kw)
        assert (mode & 1) == (0 if persist.get("DOAPP_PERSIST") == "0" else 1)
        _same(st, want)
        assert np.array_equal(r["best_traj"].view(np.uint32), want["best_traj"].view(np.uint32))


def test_pipeline_planner_takes_the_exact_fallback_on_odd_angles(capi, oracle_mod, monkeypatch):
    k, m, n, iters = 2, 64, 32, 6
    kw = dict(sigma=0.05, select_mode=1)
    start = synth.START
    init = capi.plan_lerp(start, synth.GOALS[1], n)
    nz = synth.noise(iters * k, m, 1, 5, 0.05).reshape(iters, k, m, 1, 5).copy()
    nz[0, 0, 1] = 0.0                 # x == base: exact zeros for the joints that do not move
    nz[1, 0, 2, 0, :] = 1e-41         # denormal angles: outside the verified ranges
    nz[2, 1, 5, 0, 1] = -init[3, 1]   # one angle exactly 0 in the middle of a trajectory
    nz[3, 0, 6, 0, 0] = 3e5           # |angle| >= 2^17: full-range sin / cos
    nz[4, 1, 7, 0, 3] = np.float32(2.0 ** -40)
    orc = oracle_mod.Oracle(k=k, m=m, n=n, **kw)
    orc.set_grid(synth.box_grid(64, 0.02), 0.02)
    want = orc.plan(start, init, iters, nz)
    for persist in ({}, {"DOAPP_PERSIST_V": "1"}, {"DOAPP_PERSIST": "0"}):
        r, st, mode = _run(capi, monkeypatch, persist, k, m, n, iters, nz, start, init, None, **kw)
        _same(st, want)
        assert np.float32(r["best_score"]) == np.float32(want["best_score"])
