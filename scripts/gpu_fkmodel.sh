# FK-model workloads beside the reference-chain ones
set -x
mkdir -p gpurun_out/u
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for w in c2 c2u c4 c4u; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/u/bench_$w.json 2> gpurun_out/u/bench_$w.err || tail -5 gpurun_out/u/bench_$w.err
done
python - <<'PY'
import json
for w in ("c2","c2u","c4","c4u"):
    try:
        j=json.load(open(f"gpurun_out/u/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], j["roofline"]["kernel"], "gathers/s=%.4g"%j["gather_roofline"]["achieved_gathers_per_s"])
    except Exception as e: print(w, "ERR", e)
PY
