# N-GPU round: the multi-GPU parity check and the scaling bench lines (peer mailboxes vs NCCL)
N=${1:-2}
mkdir -p gpurun_out/m
nvidia-smi -L | head -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multigpu_check.py 2>&1 | grep -v "^\[W\|^W1\|Warning" | tail -12
for w in c2 c4; do
  for cfg in 1 ${N} ${N}nccl; do
    n=${cfg%nccl}
    if [ $n = 1 ]; then
      timeout 600 python bench.py --workload $w --no-cpu --steps 10 > gpurun_out/m/bench_${w}_n$cfg.json 2> gpurun_out/m/bench_${w}_n$cfg.err
    else
      [ $cfg != $n ] && export DOAPP_NO_PEER=1 || unset DOAPP_NO_PEER
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --workload $w --no-cpu --steps 10 > gpurun_out/m/bench_${w}_n$cfg.json 2> gpurun_out/m/bench_${w}_n$cfg.err || tail -20 gpurun_out/m/bench_${w}_n$cfg.err
      unset DOAPP_NO_PEER
    fi
    python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/m/bench_${w}_n$cfg.json").read().strip().splitlines()[-1]); print("$w n=$cfg value=%.4g e2e=%.4g ms/step=%.4g it/s=%.0f"%(j["value"], j["e2e"]["value"], j["ms_per_step"], j["iterations_per_sec"]), j["config"]["parallelism"][-60:])
except Exception as e: print("$w n=$cfg ERR", e)
PY
  done
done
