# ncu --set full of the kernels of the two (f) rows — K6a / K6b (JSON out) and K7a + the dictionary kernels (documents in) —
# at 5 M records (200 M-row table), and the launch list of one JSON emission at full size.  One GPU.
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none -k regex:"doc_parse|json_emit|json_len|dict_resolve|dict_insert" -c 10 -o gpurun_out/prof_f_rows python bench.py --workload ingest --rows 200000000 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_f_rows.log 2>&1
ncu -i gpurun_out/prof_f_rows.ncu-rep --page raw --csv > gpurun_out/prof_f_rows_raw.csv 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/json_launches.csv python bench.py --workload json --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_json.log 2>&1
ls -la gpurun_out | grep "prof_f_rows\|json_launches"
