# Round 2, one GPU: GPU tests, the bench lines and the ncu evidence profiles/ is built from.
# Usage: gpurun -- 'bash scripts/gpu_r2_profile.sh'   (outputs under gpurun_out/)
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -8 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference.log 2>&1
timeout 600 python bench.py --workload sweep10m --steps 20 --warmup 5 > gpurun_out/bench_10m.log 2>&1
timeout 600 python bench.py --workload creator > gpurun_out/bench_creator.log 2>&1
timeout 600 python bench.py --workload creator-threads > gpurun_out/bench_threads.log 2>&1
timeout 300 python scripts/time_store_get.py > gpurun_out/store_get.log 2>&1; cat gpurun_out/store_get.log
ncu --query-metrics 2>/dev/null | grep -E "^dram__" > gpurun_out/metrics_dram.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 12 -c 30 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|place_group" -s 6 -c 2 -o gpurun_out/prof_sweep python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/prof_sweep.ncu-rep --page raw --csv > gpurun_out/prof_sweep_raw.csv 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 45 --csv --log-file gpurun_out/launches_store.csv python scripts/time_store_get.py > gpurun_out/ncu_store.log 2>&1
ls -la gpurun_out
