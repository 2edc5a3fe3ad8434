# Two GPUs: parity of the exchanges (NCCL and fused push) against the oracle, then the bench with each.
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 python -m pytest tests/test_push_gpu.py tests/test_multigpu_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
for ex in push compact; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --exchange $ex > gpurun_out/bench2_$ex.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_$ex.log
done
tail -5 gpurun_out/pytest_multi.log; tail -c 1500 gpurun_out/bench2_push.log
