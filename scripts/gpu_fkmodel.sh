# FK-model workload: bench beside plain C4, then instruction count of its cost kernel
set -x
mkdir -p gpurun_out/u
for w in c4 c4u; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/u/bench_$w.json 2> gpurun_out/u/bench_$w.err || tail -5 gpurun_out/u/bench_$w.err
done
python - <<'PY'
import json
for w in ("c4","c4u"):
    try:
        j=json.load(open(f"gpurun_out/u/bench_{w}.json"))
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%.2f"%j["roofline"]["avg_launch_us"], "gathers/s=%.4g"%j["gather_roofline"]["achieved_gathers_per_s"], "gfrac=%.3f"%j["gather_roofline"]["frac"])
    except Exception as e: print(w, "ERR", e)
PY
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread --clock-control none -k regex:cost_kernel -c 6 --csv --log-file gpurun_out/u/ncu_c4u.csv python bench.py --workload c4u --no-cpu --steps 1 --warmup 1 > gpurun_out/u/ncu_c4u.log 2>&1
tail -30 gpurun_out/u/ncu_c4u.csv | cut -c1-400
