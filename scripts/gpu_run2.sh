set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log

timeout 300 python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
timeout 300 python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|place_hits" -s 9 -c 3 -o gpurun_out/prof_sweep python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1

timeout 120 ./scripts/sweep_lab > gpurun_out/lab.log 2>&1
ls -la gpurun_out
