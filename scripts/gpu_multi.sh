# N-GPU round: GPU tests (incl. the 2-GPU parity check) and the scaling bench lines
N=${1:-2}
mkdir -p gpurun_out/m
nvidia-smi -L | head -8
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in c2 c4; do
  for n in 1 $N; do
    if [ $n = 1 ]; then
      timeout 600 python bench.py --workload $w --no-cpu --steps 10 > gpurun_out/m/bench_${w}_n$n.json 2> gpurun_out/m/bench_${w}_n$n.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --workload $w --no-cpu --steps 10 > gpurun_out/m/bench_${w}_n$n.json 2> gpurun_out/m/bench_${w}_n$n.err || tail -20 gpurun_out/m/bench_${w}_n$n.err
    fi
    python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/m/bench_${w}_n$n.json").read().strip().splitlines()[-1]); print("$w n=$n value=%.4g e2e=%.4g ms/step=%.4g it/s=%.0f"%(j["value"], j["e2e"]["value"], j["ms_per_step"], j["iterations_per_sec"]))
except Exception as e: print("$w n=$n ERR", e)
PY
  done
done
