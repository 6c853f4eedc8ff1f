set -x
mkdir -p gpurun_out/n
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 7 -c 1 -f -o gpurun_out/n/full_c4 python bench.py --workload c4 --steps 1 --warmup 3 --no-cpu > gpurun_out/n/ncu_full_c4.log 2>&1
ls -la gpurun_out/n
