# K7 parity tests (parser staging its documents in shared memory, and in global memory), then the document load timed
# both ways.  One GPU.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest_gpu.py tests/test_json_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_ing.log 2>&1; tail -3 gpurun_out/pytest_ing.log
TT_PARSE_STAGE=0 timeout 600 python -m pytest tests/test_ingest_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_ing_global.log 2>&1; tail -2 gpurun_out/pytest_ing_global.log
TT_PARSE_STAGE=4096 timeout 600 python -m pytest tests/test_ingest_gpu.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_ing_small.log 2>&1; tail -2 gpurun_out/pytest_ing_small.log
for v in "11264 5" "11264 4" "0 16"; do
  set -- $v
  TT_PARSE_STAGE=$1 TT_PARSE_CTAS=$2 timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest_s$1_c$2.log 2>&1; tail -1 gpurun_out/bench_ingest_s$1_c$2.log | cut -c280-420
done
