# quick GPU round: tests, the three bench workloads, instruction counts of the cost kernel
set -x
mkdir -p gpurun_out/q
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/q/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q/pytest_gpu.log
tail -5 gpurun_out/q/pytest_gpu.log
for w in c2 c4 c5; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/q/bench_$w.json 2> gpurun_out/q/bench_$w.err
done
python - <<'PY'
import json
for w in ("c2","c4","c5"):
    try:
        j=json.load(open(f"gpurun_out/q/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], "frac=%.4f"%j["roofline"]["frac"])
    except Exception as e: print(w, "ERR", e)
PY
for w in c4 c5; do
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:cost_kernel -s 7 -c 1 --csv --log-file gpurun_out/q/ncu_$w.csv python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1
grep -v "^==" gpurun_out/q/ncu_$w.csv | awk -F'","' 'NR>1{print $(NF-2), $NF}'
done
