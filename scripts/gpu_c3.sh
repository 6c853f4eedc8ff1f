timeout 300 python -m pytest tests/test_plan_gpu.py tests/test_baseline_grids_gpu.py -m gpu -q -k "replan" 2>&1 | tail -3
timeout 300 python bench.py --workload c3 --steps 10 --no-cpu 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('c3 value=%.4g e2e=%.4g ms/step=%.4g'%(j['value'], j['e2e']['value'], j['ms_per_step']))
    else: print(l[:200])
"
