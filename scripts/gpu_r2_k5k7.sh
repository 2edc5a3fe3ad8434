# K5 (prefix passes, tile size) and the K7 dictionary encoders: parity tests, then A/B timings.  One GPU.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sweep_gpu.py tests/test_ingest_gpu.py tests/test_shards_gpu.py tests/test_creator_gpu.py tests/test_point_ops_gpu.py tests/test_backlog_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_k5k7.log 2>&1; tail -5 gpurun_out/pytest_k5k7.log
for items in 16 8; do for prefix in 1 0; do
  TT_RADIX_ITEMS=$items TT_RADIX_PREFIX=$prefix timeout 300 python scripts/time_store_get.py > gpurun_out/store_get_i${items}_p${prefix}.log 2>&1; tail -4 gpurun_out/store_get_i${items}_p${prefix}.log
done; done
timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest.log 2>&1; tail -1 gpurun_out/bench_ingest.log | cut -c1-900
TT_DICT_SORTED=1 timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest_sorted.log 2>&1; tail -1 gpurun_out/bench_ingest_sorted.log | cut -c1-900
