# time cost-kernel build variants (scripts/micro/variants/lib_*.so) against the in-tree build
mkdir -p gpurun_out/v
for v in default $(ls scripts/micro/variants/ 2>/dev/null | sed 's/^lib_//; s/\.so$//'); do
  for w in c4 c4s; do
    if [ $v = default ]; then unset DOAPP_LIB; else export DOAPP_LIB=$PWD/scripts/micro/variants/lib_$v.so; fi
    timeout 300 python bench.py --workload $w --no-cpu --no-also --steps 5 > gpurun_out/v/${v}_$w.json 2> gpurun_out/v/${v}_$w.err
    python - <<PY
import json
try:
    j=json.load(open("gpurun_out/v/${v}_$w.json")); print("$v $w value=%.4g launch_us=%.2f"%(j["value"], j["roofline"]["avg_launch_us"]))
except Exception as e: print("$v $w ERR", e)
PY
  done
done
