set -x
mkdir -p gpurun_out
for mode in exact fixed; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --exchange-mode $mode > gpurun_out/bench2_$mode.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_$mode.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --exchange none > gpurun_out/bench2_none.log 2>&1
ls -la gpurun_out
