set -x
mkdir -p gpurun_out
timeout 600 python bench.py --workload mixed --steps 5 --warmup 3 > gpurun_out/mixed_n1.log 2>&1; echo "rc=$?" >> gpurun_out/mixed_n1.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --workload mixed --steps 5 --warmup 3 > gpurun_out/mixed_n2.log 2>&1; echo "rc=$?" >> gpurun_out/mixed_n2.log
ls -la gpurun_out
