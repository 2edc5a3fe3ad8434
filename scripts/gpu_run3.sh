set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_creator_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
timeout 600 python bench.py --workload creator > gpurun_out/bench_creator.log 2>&1; echo "rc=$?" >> gpurun_out/bench_creator.log

timeout 600 python bench.py --workload mixed --steps 10 --warmup 3 > gpurun_out/bench_mixed.log 2>&1; echo "rc=$?" >> gpurun_out/bench_mixed.log
ls -la gpurun_out
