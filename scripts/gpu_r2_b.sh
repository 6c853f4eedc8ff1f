# round 2, 1-GPU evidence call: GPU tests (incl. the BASELINE-grid parity cases), smoke, the new
# default bench line (C4-1M + also{c2,c5,c3}), the reference arm, the stage-2 pin dump of the
# reference's own device code, and the reference kernel's timing.
O=gpurun_out/r2b
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
timeout 300 ./oracle/_ref/ref_fk_harness $O/ref_fk_outputs.json > $O/ref_fk.log 2>&1; echo "ref_fk rc=$?"
timeout 300 ./oracle/_ref/ref_harness $O/ref_outputs.json > $O/ref.log 2>&1; echo "ref rc=$?"
timeout 300 ./oracle/_ref/ref_harness --time 20 2>&1 | grep '^{' > $O/ref_time.json; cat $O/ref_time.json
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -3 $O/bench.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref bench rc=$?"
python - <<'PY'
import json
try:
    j=json.load(open("gpurun_out/r2b/bench.json"))
    r=j["roofline"]
    print("c4s value=%.4g e2e=%.4g ms/step=%.4g launch_us=%.1f issue_frac=%.3f"%(j["value"], j["e2e"]["value"], j["ms_per_step"], r["avg_launch_us"], r.get("issue",{}).get("frac",-1)))
    for k,v in j.get("also",{}).items():
        print(k, v if isinstance(v,str) else "value=%.4g us/iter=%s e2e=%.4g"%(v["value"], v.get("us_per_iteration"), v["e2e"]["value"]))
    print("cpu", j.get("cpu_baseline"))
    jr=json.load(open("gpurun_out/r2b/bench_ref.json")); print("ref value=%.4g"%jr["value"], jr.get("reference_gpu_kernel"))
except Exception as e: print("ERR", e)
PY
