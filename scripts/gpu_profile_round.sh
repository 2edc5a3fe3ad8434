# Produces everything profiles/ is built from.  One GPU.  Usage: bash scripts/gpu_profile_round.sh
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 200 > gpurun_out/clocks.csv &
SMI=$!
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
kill $SMI
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference.log 2>&1
timeout 600 python bench.py --workload creator > gpurun_out/bench_creator.log 2>&1
timeout 600 python bench.py --workload sweep10m --steps 20 --warmup 5 > gpurun_out/bench_10m.log 2>&1
timeout 600 python bench.py --workload mixed --steps 10 --warmup 3 > gpurun_out/bench_mixed.log 2>&1
timeout 600 python bench.py --workload json --steps 5 --warmup 2 > gpurun_out/bench_json.log 2>&1
timeout 600 python bench.py --workload ingest --steps 3 --warmup 1 > gpurun_out/bench_ingest.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|place_group" -s 9 -c 2 -o gpurun_out/prof_sweep python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
