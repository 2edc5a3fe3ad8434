set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ingest_gpu.py tests/test_json_gpu.py tests/test_differential_gpu.py tests/test_point_ops_gpu.py tests/test_sweep_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_ing.log 2>&1; tail -15 gpurun_out/pytest_ing.log
timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest.log 2>&1; tail -1 gpurun_out/bench_ingest.log | cut -c1-700
timeout 300 python scripts/time_store_get.py > gpurun_out/store_get.log 2>&1; cat gpurun_out/store_get.log
