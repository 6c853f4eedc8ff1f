# what the driver runs at round end, in its order: GPU tests, smoke, reference arm, our arm
O=gpurun_out/final
mkdir -p $O
T0=$SECONDS; timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? wall $((SECONDS-T0)) s"; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
T0=$SECONDS; timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$? wall $((SECONDS-T0)) s"; tail -1 $O/bench_ref.err
T0=$SECONDS; timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$? wall $((SECONDS-T0)) s"; tail -1 $O/bench.err
python - <<'PY'
import json
j=json.load(open("gpurun_out/final/bench.json")); r=j["roofline"]; jr=json.load(open("gpurun_out/final/bench_ref.json"))
print("ours  value=%.4g e2e=%.4g ms/step=%.4g launches=%d clocks=%s"%(j["value"], j["e2e"]["value"], j["ms_per_step"], j["gpu_launches"], j["clocks"]))
print("      roofline frac=%.4f achieved=%.1f GB/s traffic=%s issue=%.3f (%.1f inst/wps) share=%.3f"%(r["frac"], r["achieved"], r["traffic"], r["issue"]["frac"], r["issue"]["thread_inst_per_wps"], r["kernel_share_of_step"]))
for k,v in j.get("also",{}).items():
    print("      also",k, v if isinstance(v,str) else "value=%.4g us/iter=%s e2e=%.4g stats=%s"%(v["value"], v.get("us_per_iteration"), v["e2e"]["value"], v.get("planner_stats")))
print("      cpu_baseline %.4g (%d cores), single thread %.4g"%(j["cpu_baseline"]["value"], j["cpu_baseline"]["cores"], j["cpu_baseline"]["single_thread"]["value"]))
print("ref   value=%.4g ms/step=%.4g  same_config=%s  e2e ratio=%.1f"%(jr["value"], jr["ms_per_step"], jr["config"]==j["config"], j["e2e"]["value"]/jr["value"]))
print("      reference_gpu_kernel", jr.get("reference_gpu_kernel"))
PY
