# bench lines of every workload + the N=2 multi-GPU run (if 2 GPUs are visible)
mkdir -p gpurun_out/b
for w in c2 c3 c4 c4m c5; do
  extra="--no-cpu"; [ $w = c2 ] && extra=""
  timeout 600 python bench.py --workload $w $extra > gpurun_out/b/bench_$w.json 2> gpurun_out/b/bench_$w.err || tail -5 gpurun_out/b/bench_$w.err
done
python - <<'PY'
import json
for w in ("c2","c3","c4","c4m","c5"):
    try:
        j=json.load(open(f"gpurun_out/b/bench_{w}.json"))
        r=j["roofline"]; g=j.get("gather_roofline",{})
        print(w, "value=%.4g e2e=%.4g ms/step=%.4g"%(j["value"], j["e2e"]["value"], j["ms_per_step"]),
              "launch_us=%s frac=%s"%(r.get("avg_launch_us"), r.get("frac")),
              "gather: ach=%.3g l2=%.3g hbm=%.3g frac=%.3f"%(g.get("achieved_gathers_per_s",0), g.get("peak_l2_gathers_per_s",0), g.get("peak_hbm_gathers_per_s",0), g.get("frac",0)) if g else "")
    except Exception as e: print(w, "ERR", e)
PY
