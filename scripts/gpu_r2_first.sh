# round 2, call 1: GPU tests at HEAD on a 2-GPU box (the 2-GPU parity test runs instead of skipping)
O=gpurun_out/r2a
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/multigpu_check.py > $O/multigpu_check_n2.log 2>&1; echo "multigpu rc=$?"
grep "multigpu_check ok\|Error\|assert" $O/multigpu_check_n2.log | tail
