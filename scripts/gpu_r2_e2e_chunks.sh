# e2e (tt_tick_host) against the number of pipeline chunks.  One GPU.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_shards_gpu.py -m gpu -x -q --timeout 500 2>&1 | tail -2
for C in 8 16 32 64; do
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-chunks $C > gpurun_out/e2e_c$C.log 2>&1
grep '^{' gpurun_out/e2e_c$C.log | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print($C, d['e2e']['value'], d['e2e']['sequential']['value'], d['ms_per_step'])"
done
