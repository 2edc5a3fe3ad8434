set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_shards_gpu.py tests/test_configs_gpu.py tests/test_push_gpu.py -m gpu -x -q --timeout 900 > gpurun_out/pytest_sh.log 2>&1; tail -25 gpurun_out/pytest_sh.log
