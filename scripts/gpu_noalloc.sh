# the no-alloc entry point: alone in a fresh process, then after every other GPU test
O=gpurun_out/noalloc
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/alone.log 2>&1; echo "alone rc=$?"; tail -3 $O/alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/all.log 2>&1; echo "all rc=$?"; tail -4 $O/all.log
