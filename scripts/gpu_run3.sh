set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sweep_gpu.py tests/test_creator_gpu.py -m gpu -q --timeout 900 -k "not full_size" > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
timeout 600 python scripts/time_store_get.py 1000000000 > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"radix|order_keys|emit_rows|minmax" -c 40 --csv --log-file gpurun_out/launches_sort.csv python scripts/time_store_get.py 1000000000 > gpurun_out/ncu.log 2>&1
ls -la gpurun_out
