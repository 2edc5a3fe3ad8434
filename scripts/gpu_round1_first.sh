set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; free -g >> gpurun_out/gpu.txt
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 -x --deselect tests/test_sweep_gpu.py::test_full_size_properties_1b > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
timeout 600 python -m pytest tests -m gpu -q --timeout 600 -k full_size > gpurun_out/pytest_1b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_1b.log
timeout 300 python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
timeout 300 python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 3 -c 2 -o gpurun_out/prof_sweep python bench.py --steps 3 --warmup 3 --rows 268435456 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
