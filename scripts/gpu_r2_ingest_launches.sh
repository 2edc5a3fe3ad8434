# Launch list of one document load (K7) and one JSON emission (K6).  One GPU.
set -x
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ingest_launches.csv python bench.py --workload ingest --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_ingest.log 2>&1
tail -2 gpurun_out/ncu_ingest.log | cut -c1-300
