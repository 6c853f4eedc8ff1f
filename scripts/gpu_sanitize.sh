# compute-sanitizer memcheck + racecheck on the small config: cost_eval, the
# three planner paths (launch per iteration, first / second generation persistent), scatter,
# distance field.  Logs -> gpurun_out/sanitize/
O=gpurun_out/sanitize
mkdir -p $O
cat > /tmp/san.py <<'PY'
import os, sys
sys.path.insert(0, '.')
import numpy as np
from doapp_b200 import capi
from tests import synth
g = synth.box_grid(64, 0.02)
for env in ({"DOAPP_PERSIST": "0"}, {"DOAPP_PERSIST_V": "1"}, {}):
    for k in ("DOAPP_PERSIST", "DOAPP_PERSIST_V"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ctx = capi.Context(k=2, m=64, n=32, sigma=0.1, select_mode=1)
    ctx.set_grid(g, 0.02)
    base = np.stack([capi.plan_lerp(synth.START, synth.GOALS[c], 32) for c in range(2)])
    ctx.cost_eval(base, synth.START, None, iteration=1)
    r = ctx.plan(synth.START, synth.GOALS[1], base[1], 12)
    print(env, ctx.plan_mode(), r["best_score"], flush=True)
    ctx.scatter(synth.cloud(5000, 0), 1, True)
    ctx.grid_distance_field()
    ctx.close()
PY
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python /tmp/san.py > $O/$tool.log 2>&1; echo "$tool rc=$?"
  grep -c "ERROR SUMMARY\|RACECHECK SUMMARY" $O/$tool.log; tail -4 $O/$tool.log
done
