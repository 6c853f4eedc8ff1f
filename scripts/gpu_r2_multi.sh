# N-GPU: parity of the sharded planner (peer mailboxes in the persistent planner, NCCL per
# iteration) with the 1-GPU plan and the oracle, then the default bench at N and C2 at N
N=${1:-2}
O=gpurun_out/r2m$N
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multigpu_check.py > $O/multigpu_check.log 2>&1; echo "multigpu_check rc=$?"
grep "multigpu_check ok\|Error\|error\|assert" $O/multigpu_check.log | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench rc=$?"; tail -3 $O/bench_n$N.err | cut -c1-300
DOAPP_NO_PEER=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 10 --workload c2 > $O/bench_c2_nccl_n$N.json 2> $O/bench_c2_nccl_n$N.err; echo "bench c2 nccl rc=$?"
DOAPP_XCHG=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --no-also > $O/bench_nccl_n$N.json 2> $O/bench_nccl_n$N.err; echo "bench c4s nccl rc=$?"
python - <<PY
import json
for f in ("bench_n$N","bench_nccl_n$N","bench_c2_nccl_n$N"):
    try:
        j=json.loads(open("$O/"+f+".json").read().strip().splitlines()[-1])
        print(f, "value=%.4g e2e=%.4g ms/step=%.4g parity=%s"%(j["value"], j["e2e"]["value"], j["ms_per_step"], j.get("parity_vs_1gpu")), j["run"]["parallelism"][-70:])
        for k,v in j.get("also",{}).items():
            print("  also",k, v if isinstance(v,str) else "value=%.4g us/iter=%s parity=%s %s"%(v["value"], v.get("us_per_iteration"), v.get("parity_vs_1gpu"), v.get("parallelism","")[-60:]))
    except Exception as e: print(f,"ERR",e)
PY
