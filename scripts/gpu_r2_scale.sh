# N-GPU, trimmed: parity of every merge path, then the default bench line (c4s + also.c2)
N=${1:-8}
O=gpurun_out/r2s$N
mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multigpu_check.py > $O/multigpu_check.log 2>&1; echo "multigpu_check rc=$?"
grep -c "multigpu_check ok" $O/multigpu_check.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench rc=$?"; grep -v "^\*\|OMP_NUM\|^W1" $O/bench_n$N.err | tail -3 | cut -c1-300
python - <<PY
import json
j=json.loads(open("$O/bench_n$N.json").read().strip().splitlines()[-1])
print("c4s n=$N value=%.4g e2e=%.4g ms/step=%.4g parity=%s"%(j["value"], j["e2e"]["value"], j["ms_per_step"], j.get("parity_vs_1gpu")), j["run"]["parallelism"][-70:])
for k,v in j.get("also",{}).items():
    print("  also",k, v if isinstance(v,str) else "value=%.4g us/iter=%s parity=%s %s"%(v["value"], v.get("us_per_iteration"), v.get("parity_vs_1gpu"), v.get("parallelism","")[-60:]))
PY
