# round 2 profiles: launch lists (gpu__time_duration) of the default bench, C2 and C3, then one
# `ncu --set full` capture of the dominant kernel of c4s, c2, c5 and of the EDT / scatter kernels.
# Each ncu command runs only after the same command has exited 0 without ncu.
O=gpurun_out/r2ncu
mkdir -p $O
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-also > $O/bench_c4s_short.json 2> $O/bench_c4s_short.err; echo "bench c4s rc=$?"
timeout 600 python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu > $O/bench_c2_short.json 2> $O/bench_c2_short.err; echo "bench c2 rc=$?"
timeout 600 python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu > $O/bench_c5_short.json 2> $O/bench_c5_short.err; echo "bench c5 rc=$?"
timeout 600 python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu > $O/bench_c3_short.json 2> $O/bench_c3_short.err; echo "bench c3 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4s.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-also > $O/ncu_launch_c4s.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_c2.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu > $O/ncu_launch_c3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cost_kernel_pk -s 30 -c 1 -f -o $O/full_c4s python bench.py --steps 1 --warmup 3 --no-cpu --no-also > $O/ncu_full_c4s.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plan_pipeline -s 3 -c 1 -f -o $O/full_c2_pipeline python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cost_kernel_pk -s 10 -c 1 -f -o $O/full_c5 python bench.py --workload c5 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c5.log 2>&1
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
timeout 300 python /tmp/edt.py; echo "edt rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"edt_|scatter_kernel" -s 4 -c 4 -f -o $O/full_aux python /tmp/edt.py > $O/ncu_full_aux.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_edt.csv python /tmp/edt.py > /dev/null 2>&1
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $b.src.csv $r
done
ls -la $O | head -40
