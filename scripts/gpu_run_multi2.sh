set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu_gpu.py tests/test_sweep_gpu.py -m gpu -q --timeout 600 -k "nccl or compact" > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
for ex in compact rows; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --exchange $ex > gpurun_out/bench2_$ex.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_$ex.log
done
ls -la gpurun_out
