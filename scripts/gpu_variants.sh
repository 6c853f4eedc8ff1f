# packed vs scalar cost kernel on C4 (131072 samples/GPU), C4m (1M samples) and C5
mkdir -p gpurun_out/v
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for v in default scalar; do
  for w in c4 c4m c5; do
    if [ $v = default ]; then unset DOAPP_PK; else export DOAPP_PK=0; fi
    timeout 300 python bench.py --workload $w --no-cpu --steps 5 > gpurun_out/v/${v}_$w.json 2> gpurun_out/v/${v}_$w.err
    python - <<PY
import json
try:
    j=json.load(open("gpurun_out/v/${v}_$w.json")); print("$v $w value=%.4g launch_us=%.2f"%(j["value"], j["roofline"]["avg_launch_us"]))
except Exception as e: print("$v $w ERR", e)
PY
  done
done
unset DOAPP_PK
for w in c4; do
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:cost_kernel -s 7 -c 1 --csv --log-file gpurun_out/v/ncu_$w.csv python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1
grep -v "^==" gpurun_out/v/ncu_$w.csv | awk -F'","' 'NR>1{print $(NF-2), $NF}'
done
