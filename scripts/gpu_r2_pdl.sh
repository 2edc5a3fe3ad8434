# Programmatic dependent launch of the tick's kernels (A/B against TT_NO_PDL=1), the document load after its
# allocations moved up front, and K7a at 9 / 12 / 16 CTAs per SM.  One GPU.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --no-extra > gpurun_out/bench_pdl.log 2>&1; tail -1 gpurun_out/bench_pdl.log | cut -c1-330
TT_NO_PDL=1 timeout 600 python bench.py --no-extra --no-cpu-baseline > gpurun_out/bench_nopdl.log 2>&1; tail -1 gpurun_out/bench_nopdl.log | cut -c1-330
timeout 300 python bench.py --workload sweep10m > gpurun_out/bench_10m.log 2>&1; tail -1 gpurun_out/bench_10m.log | cut -c1-500
TT_NO_PDL=1 timeout 300 python bench.py --workload sweep10m > gpurun_out/bench_10m_nopdl.log 2>&1; tail -1 gpurun_out/bench_10m_nopdl.log | cut -c1-500
for c in 9 12 16; do
  TT_PARSE_CTAS=$c timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest_c$c.log 2>&1; tail -1 gpurun_out/bench_ingest_c$c.log | cut -c300-520
done
timeout 300 python scripts/time_store_get.py > gpurun_out/store_get.log 2>&1; tail -3 gpurun_out/store_get.log
