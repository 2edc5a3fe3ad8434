set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_json_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_json.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_json.log
timeout 900 python bench.py --workload json --steps 5 --warmup 2 > gpurun_out/bench_json.log 2>&1; echo "rc=$?" >> gpurun_out/bench_json.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"json_emit|json_len" -c 2 -o gpurun_out/prof_json python bench.py --workload json --steps 1 --warmup 0 --no-cpu-baseline --rows 200000000 > gpurun_out/ncu_json.log 2>&1
tail -3 gpurun_out/pytest_json.log
tail -c 1500 gpurun_out/bench_json.log
