# Round 2, N GPUs of one box: the torchrun parity check and bench.py --gpus N for N = $1 (and each further argument).
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q --timeout 500 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -4 gpurun_out/pytest_multi.log
for N in "$@"; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$N.log
grep '^{' gpurun_out/bench_n$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    print({k:d.get(k) for k in ['n_gpus','ms_per_step','value','ids_at_root_ms_per_step','strong_ms_per_step','strong_tasks_per_sec','strong_collect_ms','mixed_ms']}, d['exchange']['collect_ms'], d['exchange']['ms'], d['roofline']['kernel_ms'], d['roofline']['place_kernels']['ms'], (d.get('e2e') or {}).get('value'))
"
tail -2 gpurun_out/bench_n$N.log | cut -c1-200
done
