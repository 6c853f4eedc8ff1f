# validation of the pk-kernel fallback move and the exhaustive EDT lines kernel: full GPU suite,
# EDT launch times, bench line
O=gpurun_out/r2e
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
bash scripts/gpu_edt.sh 2>&1 | tail -8
timeout 600 python bench.py --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -2 $O/bench.err
python - <<'PY'
import json
j=json.load(open("gpurun_out/r2e/bench.json")); r=j["roofline"]
print("c4s value=%.4g e2e=%.4g ms/step=%.4g launch_us=%.1f issue_frac=%.3f"%(j["value"], j["e2e"]["value"], j["ms_per_step"], r["avg_launch_us"], r.get("issue",{}).get("frac",-1)))
for k,v in j.get("also",{}).items():
    print(k, v if isinstance(v,str) else "value=%.4g us/iter=%s e2e=%.4g"%(v["value"], v.get("us_per_iteration"), v["e2e"]["value"]))
PY
