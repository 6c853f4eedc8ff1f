# v6 evidence round (HEAD after the FK model / distance field / packed-parts work):
# tests, smoke, bench lines of every workload + reference arm, launch lists, one
# `ncu --set full` capture of each kernel that changed since v5.  Output: gpurun_out/r6/
set -x
O=gpurun_out/r6
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 600 python bench.py > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench rc=$?"
for w in c2u c3 c4 c4u c4m c5; do
timeout 600 python bench.py --workload $w --no-cpu > $O/bench_$w.json 2> $O/bench_$w.err || tail -5 $O/bench_$w.err
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python - <<'PY'
import json
for w in ("c2","c2u","c3","c4","c4u","c4m","c5","ref"):
    try:
        j=json.load(open(f"gpurun_out/r6/bench_{w}.json"))
        r=j.get("roofline") or {}
        print(w, "value=%.4g"%j["value"], "e2e=%.4g"%j["e2e"]["value"], "ms/step=%.4g"%j["ms_per_step"], "launch_us=%s"%r.get("avg_launch_us"), r.get("kernel"), "frac=%s"%r.get("frac"))
    except Exception as e: print(w, "ERR", e)
PY
# launch lists
for w in c2 c2u c4u c5; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_$w.log 2>&1
done
# full captures of the kernels new since v5
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plan_persistent -s 3 -c 1 -f -o $O/full_c2u_persistent python bench.py --workload c2u --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c2u.log 2>&1
for w in c4u c5; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 7 -c 1 -f -o $O/full_$w python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > $O/ncu_full_$w.log 2>&1
done
timeout 600 ncu --set full --clock-control none -k regex:"edt_rows|edt_lines" -c 6 -f -o $O/full_edt python -m pytest tests/test_distance_field.py -m gpu -x -q > $O/ncu_full_edt.log 2>&1
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $b.src.csv $r
done
ls -la $O
