# K7 parity tests (fast + general parser, general only, small stage), then the document load timed.  One GPU.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest_gpu.py tests/test_json_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_ing.log 2>&1; tail -3 gpurun_out/pytest_ing.log
TT_PARSE_FAST=0 timeout 600 python -m pytest tests/test_ingest_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_ing_general.log 2>&1; tail -2 gpurun_out/pytest_ing_general.log
for v in 1 0; do
  TT_PARSE_FAST=$v timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest_fast$v.log 2>&1; tail -1 gpurun_out/bench_ingest_fast$v.log | cut -c280-420
done

