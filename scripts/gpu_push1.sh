# One GPU: tests of the fused placement/push kernel, then K3 variants timed in the headline bench.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_push_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_push.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_push.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_hits.log 2>&1
TT_PLACE_KERNEL=group timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_group.log 2>&1
tail -5 gpurun_out/pytest_push.log
