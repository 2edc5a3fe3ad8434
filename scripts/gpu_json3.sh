set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_json_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_json.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_json.log
timeout 900 python bench.py --workload json --steps 5 --warmup 2 > gpurun_out/bench_json.log 2>&1; echo "rc=$?" >> gpurun_out/bench_json.log
tail -3 gpurun_out/pytest_json.log
tail -c 1500 gpurun_out/bench_json.log
