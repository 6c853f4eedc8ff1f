# pk kernel change: parity tests that exercise it, then the bench lines of c5, c4s, c4
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_baseline_grids_gpu.py tests/test_plan_gpu.py tests/test_distance_field.py tests/test_weighted_update.py -m gpu -x -q 2>&1 | tail -3
for w in c5 c4s c4; do
  timeout 300 python bench.py --workload $w --no-cpu --no-also --steps 10 > /tmp/b_$w.json 2>/tmp/b_$w.err
  python - <<PY
import json
try:
    j=json.load(open("/tmp/b_$w.json")); print("$w value=%.4g launch_us=%.2f"%(j["value"], j["roofline"]["avg_launch_us"]))
except Exception as e: print("$w ERR", e, open("/tmp/b_$w.err").read()[-300:])
PY
done
