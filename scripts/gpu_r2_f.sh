# final validation of the round: GPU suite, smoke, ncu re-capture of the changed pk kernel (c4s, c5),
# default bench + reference arm
O=gpurun_out/r2f
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -3 $O/pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-also > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:cost_kernel_pk -s 30 -c 1 -f -o $O/full_c4s python bench.py --steps 1 --warmup 3 --no-cpu --no-also > $O/ncu_full_c4s.log 2>&1
timeout 600 python bench.py --workload c5 --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:cost_kernel_pk -s 10 -c 1 -f -o $O/full_c5 python bench.py --workload c5 --steps 1 --warmup 3 --no-cpu > $O/ncu_full_c5.log 2>&1
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $b.src.csv $r
done
ls $O
