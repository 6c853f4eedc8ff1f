# round 2: full GPU test suite (stress, host node, BASELINE grids), sanitizer logs, bench lines
O=gpurun_out/r2d
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -2 $O/bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
j=json.load(open("gpurun_out/r2d/bench.json")); r=j["roofline"]
print("c4s value=%.4g e2e=%.4g ms/step=%.4g launch_us=%.1f issue_frac=%.3f"%(j["value"], j["e2e"]["value"], j["ms_per_step"], r["avg_launch_us"], r.get("issue",{}).get("frac",-1)))
for k,v in j.get("also",{}).items():
    print(k, v if isinstance(v,str) else "value=%.4g us/iter=%s e2e=%.4g"%(v["value"], v.get("us_per_iteration"), v["e2e"]["value"]))
PY
bash scripts/gpu_sanitize.sh
