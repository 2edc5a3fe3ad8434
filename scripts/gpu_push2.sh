# Two GPUs: parity of the exchanges (NCCL and fused push) against the oracle, then the bench with each.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_push_gpu.py tests/test_multigpu_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
for v in 8 1 2 4; do
TT_PUSH_VARIANT=$v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --exchange push > gpurun_out/bench2_push_v$v.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_push_v$v.log
done
tail -5 gpurun_out/pytest_multi.log
