O=gpurun_out/p2t
mkdir -p $O
DOAPP_P2_TRACE=1 timeout 300 python - > $O/trace.log 2>&1 <<'PY'
import os, sys
sys.path.insert(0, '.')
import numpy as np
from doapp_b200 import capi
from tests import synth
g = synth.box_grid(128, 0.01)
for depth in ("1", "6", "8", "10"):
    os.environ["DOAPP_PERSIST_DEPTH"] = depth
    ctx = capi.Context(k=1, m=1024, n=64, sigma=0.05, select_mode=1)
    ctx.set_grid(g, 0.01)
    init = capi.plan_lerp(synth.START, synth.GOALS[1], 64)
    ctx.plan(synth.START, synth.GOALS[1], init, 100)
    for _ in range(3):
        ctx.plan_bench(100, 1)
    ms = min(ctx.plan_bench(100, 1) for _ in range(10))
    print("depth", depth, "us/iter", ms * 10, flush=True)
    sys.stderr.flush()
    ctx.plan_stats()
    ctx.close()
PY
cat $O/trace.log
