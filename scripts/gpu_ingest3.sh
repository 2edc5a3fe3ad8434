set -x
mkdir -p gpurun_out
python scripts/all_kernels_small.py > gpurun_out/all_small.log 2>&1; echo "rc=$?" >> gpurun_out/all_small.log
timeout 900 python -m pytest tests/test_ingest_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_ingest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ingest.log
timeout 900 python bench.py --workload ingest --steps 3 --warmup 1 > gpurun_out/bench_ingest.log 2>&1; echo "rc=$?" >> gpurun_out/bench_ingest.log
tail -4 gpurun_out/all_small.log; tail -15 gpurun_out/pytest_ingest.log; tail -c 1500 gpurun_out/bench_ingest.log
