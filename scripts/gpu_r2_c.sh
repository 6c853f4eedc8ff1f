# round 2: first run of the second-generation persistent planner + the full GPU test suite
O=gpurun_out/r2c
mkdir -p $O
timeout 300 python scripts/p2_check.py > $O/p2_check.log 2>&1; echo "p2_check rc=$?"; tail -32 $O/p2_check.log
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
