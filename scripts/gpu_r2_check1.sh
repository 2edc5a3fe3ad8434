# One-GPU check: the whole GPU suite, smoke, the headline bench line, config 2, the document load, the ordered list.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/pytest.log 2>&1; tail -3 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log | cut -c1-330
timeout 300 python bench.py --workload sweep10m > gpurun_out/bench_10m.log 2>&1; tail -1 gpurun_out/bench_10m.log | cut -c1-400
timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ingest.log 2>&1; tail -1 gpurun_out/bench_ingest.log | cut -c280-420
timeout 300 python scripts/time_store_get.py > gpurun_out/store_get.log 2>&1; tail -2 gpurun_out/store_get.log
