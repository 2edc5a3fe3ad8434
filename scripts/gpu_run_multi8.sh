# One 8-GPU box: parity at 4 ranks, then the headline bench at 2/4/8 GPUs (fused push exchange, the
# default), and at 8 GPUs the NCCL compact exchange, no exchange, strong scaling and the config-5 mix.
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multigpu_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
run() { # name, nproc, extra args
  name=$1; n=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 10 --warmup 3 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$?" >> gpurun_out/$name.log
}
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.log 2>&1
for n in 2 4 8; do
  if [ $n -le $N ]; then run bench_n$n $n; fi
done
TT_PUSH_VARIANT=8 run bench_push8x4_n$N $N --no-e2e
run bench_compact_n$N $N --no-e2e --exchange compact
run bench_none_n$N $N --no-e2e --exchange none
run bench_strong_n$N $N --no-e2e --scaling strong
run bench_mixed_n$N $N --workload mixed
ls -la gpurun_out
