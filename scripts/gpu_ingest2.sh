set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ingest_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_ingest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ingest.log
timeout 900 python bench.py --workload ingest --steps 3 --warmup 1 > gpurun_out/bench_ingest.log 2>&1; echo "rc=$?" >> gpurun_out/bench_ingest.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"doc_parse" -c 1 -o gpurun_out/prof_ingest python bench.py --workload ingest --steps 1 --warmup 0 --no-cpu-baseline --rows 200000000 > gpurun_out/ncu_ingest.log 2>&1
tail -15 gpurun_out/pytest_ingest.log
tail -c 1800 gpurun_out/bench_ingest.log
