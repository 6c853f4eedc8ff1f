# ncu --set full capture of the persistent planner on C2 (bench --workload c2), raw + source pages
O=gpurun_out/ncu_c2
mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plan_pipeline -s 3 -c 1 -f -o $O/full_c2_pipeline python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu --no-also > $O/ncu_full_c2.log 2>&1
tail -3 $O/ncu_full_c2.log
for r in $O/*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $r --page source --csv --print-source sass > $b.src.csv 2>/dev/null
  python scripts/sass_mix.py $b.src.csv > $b.sass_mix.txt 2>/dev/null
  rm -f $r
done
ls -la $O
