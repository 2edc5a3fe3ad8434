set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ingest_gpu.py tests/test_json_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_ingest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ingest.log
tail -40 gpurun_out/pytest_ingest.log
