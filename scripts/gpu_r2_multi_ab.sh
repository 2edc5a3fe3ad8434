# 8 GPUs: parity check, then the weak-scaled headline with the one-byte and the two-byte exchange forms, then the full line.
set -x
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q --timeout 500 > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -4 gpurun_out/pytest_multi.log
for X in rows rows16; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-extra --exchange $X > gpurun_out/ab_${N}_$X.log 2>&1
grep '^{' gpurun_out/ab_${N}_$X.log | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('N=$N $X', round(d['ms_per_step'],4), '%.3e'%d['value'], 'collect', round(d['exchange']['collect_ms'],3), 'outside', round(d['exchange']['ms'],3), 'k1', round(d['roofline']['kernel_ms'],3), 'k3', round(d['roofline']['place_kernels']['ms'],3))"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1
grep '^{' gpurun_out/bench_n$N.log | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('N=$N full', {k:d.get(k) for k in ['ms_per_step','value','ids_at_root_ms_per_step','strong_ms_per_step','strong_tasks_per_sec','strong_collect_ms','mixed_ms']}, d['exchange']['collect_ms'], d['exchange']['ms'])"
