# packed kernel with 2 / 4 parts per sample: tests, then C5 / C4 / c4u bench lines
set -x
mkdir -p gpurun_out/p
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/p/pytest_gpu.log 2>&1; tail -6 gpurun_out/p/pytest_gpu.log
for w in c5 c4 c2; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/p/bench_$w.json 2> gpurun_out/p/bench_$w.err || tail -5 gpurun_out/p/bench_$w.err
done
DOAPP_PK=0 timeout 600 python bench.py --workload c5 --no-cpu > gpurun_out/p/bench_c5_scalar.json 2> gpurun_out/p/bench_c5_scalar.err
python - <<'PY'
import json
for w in ("c5","c5_scalar","c4","c2"):
    try:
        j=json.load(open(f"gpurun_out/p/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], j["roofline"]["kernel"])
    except Exception as e: print(w, "ERR", e)
PY
