set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multigpu_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
for n in 2 4 8; do
if [ $n -le $N ]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$n.log
fi
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --scaling strong > gpurun_out/bench_strong_n$N.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --exchange ids > gpurun_out/bench_ids_n$N.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29536 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --exchange none > gpurun_out/bench_none_n$N.log 2>&1
ls -la gpurun_out
