set -x
mkdir -p gpurun_out
for v in 8 12 13 14 15 16; do
TT_PLACE_VARIANT=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_py$v.log 2>&1
done
