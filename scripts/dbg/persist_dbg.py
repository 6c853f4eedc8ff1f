import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from tests import synth
import oracle
from doapp_b200 import capi
k, m, n, d = 1, 64, 32, 5
for select in (0, 1):
  for ext in (True, False):
    for iters in (1, 2, 6):
        g = synth.box_grid(64, 0.02)
        start, goal = synth.START[:d], synth.GOALS[0][:d]
        init = capi.plan_lerp(start, goal, n)
        nz = synth.noise(iters * k, m, 1, d, 0.05).reshape(iters, k, m, 1, d) if ext else None
        res = {}
        for mode in ("0", "1"):
            os.environ["DOAPP_PERSIST"] = mode
            ctx = capi.Context(k=k, m=m, n=n, d=d, sigma=0.05, select_mode=select)
            ctx.set_grid(g, 0.02)
            a = ctx.plan(start, goal, init, iters, nz)
            st = ctx.plan_state()
            res[mode] = (st["found_better"].copy(), st["trajectories"].copy(), st["rounded"].copy(), a["best_score"])
            ctx.close()
        orc = oracle.Oracle(k=k, m=m, n=n, d=d, sigma=0.05, select_mode=select)
        orc.set_grid(g, 0.02)
        b = orc.plan(start, init, iters, nz)
        print("select", select, "ext", ext, "iters", iters, "found", res["0"][0], res["1"][0], b["found_better"],
              "traj equal 0/1:", np.array_equal(res["0"][1], res["1"][1]), "0/orc:", np.array_equal(res["0"][1], b["trajectories"]),
              "rounded", res["0"][2], res["1"][2], b["rounded"], "score", res["0"][3], res["1"][3], b["best_score"])
