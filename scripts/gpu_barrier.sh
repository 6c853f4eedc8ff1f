# persistent planner with the software-pipelined loop (speculative evaluation behind the barrier): tests, C2 / c2u lines
set -x
mkdir -p gpurun_out/b4
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/b4/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/b4/pytest_gpu.log
for w in c2 c2u; do
timeout 300 python bench.py --workload $w --no-cpu > gpurun_out/b4/bench_$w.json 2> gpurun_out/b4/bench_$w.err || tail -5 gpurun_out/b4/bench_$w.err
done
python - <<'PY'
import json
for w in ("c2","c2u"):
    try:
        j=json.load(open(f"gpurun_out/b4/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], j["roofline"]["kernel"])
    except Exception as e: print(w, "ERR", e)
PY
