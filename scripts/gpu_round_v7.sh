# v7 evidence round (HEAD after the persistent-planner barrier / pipelining work):
# tests, smoke, default bench line + reference arm, c2u / c4 / c5 lines, launch lists of
# c2 / c2u, one `ncu --set full` capture of the persistent planner.  Output: gpurun_out/r7/
set -x
O=gpurun_out/r7
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 700 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 400 python bench.py > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench rc=$?"
for w in c2u c4 c5; do
timeout 300 python bench.py --workload $w --no-cpu > $O/bench_$w.json 2> $O/bench_$w.err || tail -5 $O/bench_$w.err
done
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python - <<'PY'
import json
for w in ("c2","c2u","c4","c5","ref"):
    try:
        j=json.load(open(f"gpurun_out/r7/bench_{w}.json"))
        r=j.get("roofline") or {}
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%s"%r.get("avg_launch_us"), r.get("kernel"), "frac=%s"%r.get("frac"))
    except Exception as e: print(w, "ERR", e)
PY
for w in c2 c2u; do
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_$w.log 2>&1
done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:plan_persistent -s 3 -c 1 -f -o $O/full_c2_persistent python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c2.log 2>&1
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $b.src.csv $r
done
ls -la $O
