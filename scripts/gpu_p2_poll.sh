# collect / send warp back-off sweep on C2 (1 GPU): plan_bench best of 20, us per iteration
cat > /tmp/poll.py <<'PY'
import os, sys
sys.path.insert(0, '.')
from doapp_b200 import capi
from tests import synth
g = synth.box_grid(128, 0.01)
ctx = capi.Context(k=1, m=1024, n=64, sigma=0.05, select_mode=1)
ctx.set_grid(g, 0.01)
init = capi.plan_lerp(synth.START, synth.GOALS[1], 64)
r = ctx.plan(synth.START, synth.GOALS[1], init, 100)
for _ in range(3):
    ctx.plan_bench(100, 1)
ms = sorted(ctx.plan_bench(100, 1) for _ in range(20))
print("poll_ns", os.environ.get("DOAPP_P2_POLL_NS"), "depth", os.environ.get("DOAPP_PERSIST_DEPTH"), "us/iter best %.3f median %.3f" % (ms[0] * 10, ms[10] * 10), ctx.plan_stats(), "score", r["best_score"], flush=True)
ctx.close()
PY
for d in 8 6; do for p in 0 50 100 200 400 800; do DOAPP_PERSIST_DEPTH=$d DOAPP_P2_POLL_NS=$p timeout 120 python /tmp/poll.py; done; done
