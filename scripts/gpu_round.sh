set -x
mkdir -p gpurun_out/r1
nvidia-smi -L > gpurun_out/r1/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1/pytest_gpu.log
tail -5 gpurun_out/r1/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r1/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/r1/bench_c2.json 2> gpurun_out/r1/bench_c2.err; echo "bench rc=$?"
timeout 600 python bench.py --workload c4 --no-cpu > gpurun_out/r1/bench_c4.json 2> gpurun_out/r1/bench_c4.err
timeout 600 python bench.py --workload c5 --no-cpu > gpurun_out/r1/bench_c5.json 2> gpurun_out/r1/bench_c5.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1/bench_ref.json 2> gpurun_out/r1/bench_ref.err
cat gpurun_out/r1/bench_c2.json gpurun_out/r1/bench_c4.json gpurun_out/r1/bench_c5.json gpurun_out/r1/bench_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1/ncu_launch.log 2>&1
for w in c2 c4 c5; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 6 -c 2 -f -o gpurun_out/r1/full_$w python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > gpurun_out/r1/ncu_full_$w.log 2>&1
done
ls -la gpurun_out/r1
