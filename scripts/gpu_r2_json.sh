# K6: parity tests, then the overdue list of the 1 B-row table as JSON with 64 / 128 / 256 records per CTA.  One GPU.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_json_gpu.py tests/test_ingest_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_json.log 2>&1; tail -2 gpurun_out/pytest_json.log
for th in 128 64 256; do
  TT_JSON_THREADS=$th timeout 600 python bench.py --workload json --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/bench_json_t$th.log 2>&1; tail -1 gpurun_out/bench_json_t$th.log | cut -c290-520
done
