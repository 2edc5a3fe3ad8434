# Source-level profile of K7a (doc_parse_kernel) at 5 M documents.  One GPU.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --workload ingest --rows 200000000 --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/ingest_small.log 2>&1; tail -1 gpurun_out/ingest_small.log | cut -c1-900
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"doc_parse_fast" -s 1 -c 1 -o gpurun_out/prof_parse python bench.py --workload ingest --rows 200000000 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_parse.log 2>&1
ncu -i gpurun_out/prof_parse.ncu-rep --page source --csv > gpurun_out/prof_parse_source.csv 2>/dev/null
ncu -i gpurun_out/prof_parse.ncu-rep --page raw --csv > gpurun_out/prof_parse_raw.csv 2>/dev/null
ls -la gpurun_out | grep prof_parse
