# Full GPU evidence round: tests, smoke, bench lines of every workload + reference arm,
# ncu launch lists and one `ncu --set full` capture per kernel.  Output: gpurun_out/r1/
set -x
O=gpurun_out/r1
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench rc=$?"
for w in c3 c4 c4m c5; do
timeout 600 python bench.py --workload $w --no-cpu > $O/bench_$w.json 2> $O/bench_$w.err
done
DOAPP_PERSIST=0 timeout 600 python bench.py --no-cpu > $O/bench_c2_graph.json 2> $O/bench_c2_graph.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
# launch lists (every kernel of the same command, cold-cache serialised)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_c2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_c4.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > $O/ncu_launch_c3.log 2>&1
# full captures: persistent planner (C2), packed cost kernel (C4, C4m), scalar T=4 cost kernel (C5),
# scatter (C3), update-vector + depth kernels (tests)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:plan_persistent -s 3 -c 1 -f -o $O/full_c2_persistent python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c2.log 2>&1
for w in c4 c5; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 7 -c 1 -f -o $O/full_$w python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > $O/ncu_full_$w.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scatter_kernel -s 20 -c 1 -f -o $O/full_c3_scatter python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c3.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"update_vector|depth_scatter" -c 4 -f -o $O/full_aux python -m pytest tests/test_weighted_update.py tests/test_depth_to_grid.py -m gpu -x -q > $O/ncu_full_aux.log 2>&1
# keep the evidence small: raw page + opcode mix as text, drop the binary reports
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $b.src.csv $r
done
ls -la $O
