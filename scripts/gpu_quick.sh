# quick GPU round: tests, bench workloads, instruction counts of the cost kernel
set -x
mkdir -p gpurun_out/q
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/q/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q/pytest_gpu.log
tail -15 gpurun_out/q/pytest_gpu.log
for w in c2 c4 c5; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/q/bench_$w.json 2> gpurun_out/q/bench_$w.err || tail -5 gpurun_out/q/bench_$w.err
done
DOAPP_PERSIST=0 timeout 600 python bench.py --workload c2 --no-cpu > gpurun_out/q/bench_c2_nopersist.json 2> gpurun_out/q/bench_c2_nopersist.err
python - <<'PY'
import json
for w in ("c2","c2_nopersist","c4","c5"):
    try:
        j=json.load(open(f"gpurun_out/q/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], "frac=%.4f"%j["roofline"]["frac"], "launches", j["gpu_launches"])
    except Exception as e: print(w, "ERR", e)
PY
