set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_json_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_json.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_json.log
tail -30 gpurun_out/pytest_json.log
