O=gpurun_out/edt
mkdir -p $O
timeout 600 python -m pytest tests/test_distance_field.py tests/test_zz_no_alloc.py tests/test_host_gpu.py -m gpu -q 2>&1 | tail -4
cat > /tmp/edt.py <<'PY'
import sys
sys.path.insert(0, '.')
import numpy as np
from doapp_b200 import capi
from tests import synth
ctx = capi.Context(k=1, m=64, n=32)
for _ in range(3):
    ctx.set_grid(np.zeros((128, 128, 128), np.uint8), 0.01)
    ctx.scatter(synth.cloud(300000, 0), 1, True)
    ctx.grid_distance_field()
ctx.close()
PY
timeout 300 python /tmp/edt.py && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_edt.csv python /tmp/edt.py > /dev/null 2>&1
grep "edt_" $O/launches_edt.csv | awk -F'","' '{print $5, $(NF-1), $NF}' | cut -c1-120
