"""Quick check of plan_pipeline_kernel (second-generation persistent planner) against the
first-generation kernel and the launch-per-iteration path: bit-identical results, timing."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from doapp_b200 import capi  # noqa: E402
from tests import synth  # noqa: E402


def run(env, k, m, n, iters, sigma, grid, res, reps=0):
    for key in ("DOAPP_PERSIST", "DOAPP_PERSIST_V", "DOAPP_PERSIST_DEPTH"):
        os.environ.pop(key, None)
    os.environ.update(env)
    ctx = capi.Context(k=k, m=m, n=n, sigma=sigma, select_mode=1)
    ctx.set_grid(grid, res)
    init = capi.plan_lerp(synth.START, synth.GOALS[1], n)
    t0 = time.perf_counter()
    r = ctx.plan(synth.START, synth.GOALS[1], init, iters)
    dt = time.perf_counter() - t0
    st = ctx.plan_state()
    mode = ctx.plan_mode()
    stats = ctx.plan_stats()
    ms = None
    if reps:
        for _ in range(3):
            ctx.plan_bench(iters, 1)
        ms = min(ctx.plan_bench(iters, 1) for _ in range(reps))
    ctx.close()
    return r, st, mode, stats, dt, ms


def main():
    g = synth.box_grid(128, 0.01)
    shapes = [(1, 64, 32, 5, 0.05), (1, 1024, 64, 100, 0.05), (3, 200, 64, 6, 0.05), (2, 1024, 64, 60, 0.2),
              (1, 512, 128, 20, 0.05), (2, 100, 33, 8, 0.1), (1, 2000, 50, 10, 0.05)]
    ok = True
    for (k, m, n, iters, sigma) in shapes:
        ref = run({"DOAPP_PERSIST": "0"}, k, m, n, iters, sigma, g, 0.01)
        for env in ({"DOAPP_PERSIST_V": "1"}, {}, {"DOAPP_PERSIST_DEPTH": "1"}, {"DOAPP_PERSIST_DEPTH": "6"}):
            got = run(env, k, m, n, iters, sigma, g, 0.01)
            same = (np.array_equal(ref[0]["best_traj"].view(np.uint32), got[0]["best_traj"].view(np.uint32))
                    and np.array_equal(ref[1]["trajectories"].view(np.uint32), got[1]["trajectories"].view(np.uint32))
                    and np.array_equal(ref[1]["scores"].view(np.uint32), got[1]["scores"].view(np.uint32))
                    and np.array_equal(ref[1]["found_better"], got[1]["found_better"]))
            ok &= same
            print(f"k={k} m={m} n={n} iters={iters} env={env} mode={got[2]} same={same} stats={got[3]} "
                  f"first_call={got[4]*1e3:.1f} ms", flush=True)
    for env in ({"DOAPP_PERSIST_V": "1"}, {"DOAPP_PERSIST_DEPTH": "1"}, {"DOAPP_PERSIST_DEPTH": "2"}, {},
                {"DOAPP_PERSIST_DEPTH": "4"}, {"DOAPP_PERSIST_DEPTH": "6"}):
        got = run(env, 1, 1024, 64, 100, 0.05, g, 0.01, reps=20)
        print(f"C2 env={env} mode={got[2]} plan_bench best {got[5]*1e3:.1f} us per plan = "
              f"{got[5]*10:.3f} us/iteration stats={got[3]}", flush=True)
    print("P2 CHECK", "OK" if ok else "FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
